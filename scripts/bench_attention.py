#!/usr/bin/env python
"""Micro-benchmark of K2 (ca_joint_attention) against the bf16 tensor roofline.
FLOPs per launch = 4*128*H*B*[(T+P)^2 + C*(C+P)].  impl: 0 = default (polynomial
exp2 on every 4th pair), 2 = all MUFU, 3 = polynomial on every 2nd pair, 1 = v0 reference kernel."""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from conceptattention_b200 import _cabi, ops  # noqa: E402

SHAPES = [  # name, B, H, P, T, C
    ("cfg2 double block 1024^2", 1, 24, 4096, 256, 4),
    ("cfg2 single block 1024^2", 1, 24, 4096, 256, 0),
    ("cfg3 512^2 batch 16", 16, 24, 1024, 256, 2),
    ("cfg4 dev 1024^2 C=20", 1, 24, 4096, 512, 20),
    ("cfg5 2048^2 C=32", 1, 24, 16384, 256, 32),
    ("cfg5 head-sharded over 8 GPUs (3 heads)", 1, 3, 16384, 256, 32),
    ("cfg5 head-sharded over 2 GPUs (12 heads)", 1, 12, 16384, 256, 32),
]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--impls", default="0,2,3")
    ap.add_argument("--only", default="")
    ap.add_argument("--iters", type=int, default=20)
    args = ap.parse_args()
    pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(
        os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}
    lib = _cabi.load()
    # CA_STREAMK=-1: never split the last wave (A/B against the default plan)
    lib.ca_debug_set_attention_streamk(int(os.environ.get("CA_STREAMK", "0")))
    for name, B, H, P, T, C in SHAPES:
        if args.only not in name:
            continue
        N = P + T + C
        sets = [[(torch.randn(B, H, N, 128, device="cuda") * s).to(torch.bfloat16) for s in (1.0, 1.0, 1.0)]
                for _ in range(2)]
        flops = 4 * 128 * H * B * ((T + P) ** 2 + C * (C + P))
        for impl in [int(i) for i in args.impls.split(",")]:
            lib.ca_debug_set_attention_impl(impl)
            for i in range(3):
                ops.joint_attention(*sets[i % 2], P, T, C)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(args.iters):
                ops.joint_attention(*sets[i % 2], P, T, C)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / args.iters
            tf = flops / ms / 1e9
            print(json.dumps({"kernel": "ca_joint_attention", "impl": impl, "shape": name, "B": B, "H": H,
                              "P": P, "T": T, "C": C, "ms": round(ms, 4), "TFLOPs": round(tf, 1),
                              "frac_burst_peak": round(tf / pk["bf16_tflops"], 3),
                              "frac_sustained_peak": round(tf / pk["bf16_tflops_sustained"], 3)}), flush=True)
        lib.ca_debug_set_attention_impl(0)
        del sets
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
