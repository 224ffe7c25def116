#!/usr/bin/env python
"""Parity of the K2 kernels named per call (ca_joint_attention_ex: 1 v3, 2 v7, 3 v11) against the
default choice and the fp32 oracle on the ATTN_CASES of tests/test_gpu_kernels.py (timing is
scripts/bench_attention.py).  usage: check_k2_v5.py IMPL [IMPL...]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conceptattention_b200 import ops  # noqa: E402
from test_gpu_kernels import ATTN_CASES, _attention_oracle, rnd  # noqa: E402

CASES = ATTN_CASES + [(1, 2, 2200, 200, 5, 0), (1, 3, 4096, 256, 4, 0), (2, 2, 1024, 256, 2, 0)]


def main():
    impls = [int(a) for a in sys.argv[1:]] or [2, 3]
    bad = 0
    for B, H, P, T, C, mode in CASES:
        N = P + T + C
        q, k, v = rnd(B, H, N, 128, seed=20, scale=1.5), rnd(B, H, N, 128, seed=21), rnd(B, H, N, 128, seed=22)
        ri, rt = _attention_oracle(q.float(), k.float(), v.float(), P, T, C, mode)
        bi, bt = ops.joint_attention(q.cuda(), k.cuda(), v.cuda(), P, T, C, mode)
        for impl in impls:
            oi, ot = ops.joint_attention(q.cuda(), k.cuda(), v.cuda(), P, T, C, mode, impl=impl)
            torch.cuda.synchronize()
            err = max((oi.float().cpu() - ri).abs().max().item() if P else 0.0,
                      (ot.float().cpu() - rt).abs().max().item() if T + C else 0.0)
            base = max((bi.float().cpu() - ri).abs().max().item() if P else 0.0,
                       (bt.float().cpu() - rt).abs().max().item() if T + C else 0.0)
            ok = err <= max(1.5 * base, 1e-2) and not (torch.isnan(oi.float()).any() or torch.isnan(ot.float()).any())
            bad += not ok
            print(f"impl {impl} B={B} H={H} P={P} T={T} C={C} mode={mode}: max abs err vs oracle {err:.3e} "
                  f"(default kernel {base:.3e}) {'ok' if ok else 'MISMATCH'}", flush=True)
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
