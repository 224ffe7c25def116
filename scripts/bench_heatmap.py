#!/usr/bin/env python
"""Micro-benchmark of the heatmap kernels (K3+K4) against the HBM roofline.

For each shape: rotate over enough distinct input buffers that every launch streams
from HBM (total > 2x the 126 MB L2), time with CUDA events, report achieved GB/s on the
ALGORITHMIC bytes 2*P*K + 2*C*K + 2*4*C*P per batch element and the
fraction of the measured copy bandwidth (MEASURED_PEAKS.json).  impl 0 = TMA/tcgen05
streaming kernel, impl 1 = SIMT kernel."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from conceptattention_b200 import _cabi, ops  # noqa: E402

SHAPES = [  # name, B, P, C, K, per_head
    ("cfg2 1024^2 C=4 output space", 1, 4096, 4, 3072, False),
    ("cfg2 1024^2 C=4 cross space (per head)", 1, 4096, 4, 3072, True),
    ("cfg3 512^2 C=2 batch 32", 32, 1024, 2, 3072, False),
    ("cfg4 1024^2 C=20", 1, 4096, 20, 3072, False),
    ("cfg5 2048^2 C=32", 1, 16384, 32, 3072, False),
    ("cfg5 2048^2 C=32 cross space", 1, 16384, 32, 3072, True),
]


def main():
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="")
    ap.add_argument("--shape", action="append", default=[], help="B,P,C,K extra shape")
    args = ap.parse_args()
    peak = 6650.0
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        peak = json.load(open(path))["hbm_gbs"]
    lib = _cabi.load()
    dev = "cuda:0"
    shapes = list(SHAPES)
    for sh in args.shape:
        B, P, C, K = map(int, sh.split(","))
        shapes.append((f"custom B={B} P={P} C={C} K={K}", B, P, C, K, False))
    for name, B, P, C, K, per_head in shapes:
        if args.only not in name:
            continue
        nbytes = B * (2 * P * K + 2 * C * K + 8 * C * P)
        nbuf = max(2, int(300e6 // (B * P * K * 2)) + 1)
        if per_head:
            H = K // 128
            bufs = [torch.randn(B, H, P + C, 128, device=dev).to(torch.bfloat16) * 0.2 for _ in range(nbuf)]
            pairs = [(b[:, :, :P], b[:, :, P:]) for b in bufs]
        else:
            pairs = [(torch.randn(B, P, K, device=dev).to(torch.bfloat16) * 0.2,
                      torch.randn(B, C, K, device=dev).to(torch.bfloat16) * 0.2) for _ in range(nbuf)]
        acc = torch.zeros(B, C, P, device=dev)
        # same-size comparator: a device copy moving the same number of bytes
        src = torch.empty(nbytes // 2, dtype=torch.uint8, device=dev)
        dsts = [torch.empty_like(src) for _ in range(nbuf)]
        for d in dsts:
            d.copy_(src)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(10 * nbuf):
            dsts[i % nbuf].copy_(src)
        e1.record()
        torch.cuda.synchronize()
        copy_us = e0.elapsed_time(e1) * 1e3 / (10 * nbuf)
        del src, dsts
        for impl, split in [(0, 0)]:        # the launcher picks kernel and split-K factor by shape
            for i in range(3 * nbuf):
                ops.heatmap_accumulate(*pairs[i % nbuf], acc, "softmax", False, per_head=per_head)
            torch.cuda.synchronize()
            iters = 10 * nbuf
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(iters):
                ops.heatmap_accumulate(*pairs[i % nbuf], acc, "softmax", False, per_head=per_head)
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / iters
            gbs = nbytes / us / 1e3
            print(json.dumps({"kernel": "ca_heatmap_accumulate", "impl": "tcgen05+TMA" if impl == 0 else "SIMT",
                              "shape": name, "B": B, "P": P, "C": C, "K": K, "per_head": per_head,
                              "algorithmic_bytes": nbytes, "us_per_launch": round(us, 2),
                              "achieved_GBps": round(gbs, 1), "peak_GBps": peak,
                              "frac": round(gbs / peak, 3), "buffers_rotated": nbuf, "split": split,
                              "same_bytes_copy_us": round(copy_us, 2)}), flush=True)
        del pairs, acc
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
