#!/usr/bin/env python
"""Experimental persistent K2 (ca_debug_set_attention_impl(14)) against the default kernel:
bit equality on shapes with one, several and many units per SM (and a forced stream-K plan),
then an A/B timing at the config-2 and config-3 attention shapes.  One JSON line per case."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from conceptattention_b200 import _cabi, ops  # noqa: E402

BF16 = torch.bfloat16
lib = _cabi.load()


def run(impl, q, k, v, P, T, C, mode=0):
    lib.ca_debug_set_attention_impl(impl)
    try:
        o = ops.joint_attention(q, k, v, P, T, C, mode)
        torch.cuda.synchronize()
    finally:
        lib.ca_debug_set_attention_impl(0)
    return o


def timed(impl, q, k, v, P, T, C, iters=20):
    lib.ca_debug_set_attention_impl(impl)
    try:
        for _ in range(3):
            ops.joint_attention(q, k, v, P, T, C)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            ops.joint_attention(q, k, v, P, T, C)
        e1.record()
        torch.cuda.synchronize()
    finally:
        lib.ca_debug_set_attention_impl(0)
    return e0.elapsed_time(e1) / iters


def main():
    g = torch.Generator().manual_seed(5)
    bad = 0
    cases = [(1, 1, 200, 100, 1, 0, 0), (1, 2, 144, 72, 3, 1, 0), (2, 3, 256, 256, 4, 0, 0),
             (1, 2, 2200, 200, 5, 0, 0), (2, 24, 1024, 256, 4, 0, 0), (1, 24, 4096, 256, 4, 0, 0),
             (1, 2, 2000, 200, 3, 0, 7), (1, 2, 2200, 200, 5, 0, 24)]
    for B, H, P, T, C, mode, sk in cases:
        N = P + T + C
        q, k, v = ((torch.randn(B, H, N, 128, generator=g) * s).to(BF16).cuda() for s in (1.5, 1.0, 1.0))
        lib.ca_debug_set_attention_streamk(sk)
        try:
            a = run(0, q, k, v, P, T, C, mode)
            b = run(14, q, k, v, P, T, C, mode)
        finally:
            lib.ca_debug_set_attention_streamk(0)
        same = torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
        bad += not same
        print(json.dumps({"case": [B, H, P, T, C, mode, sk], "bit_identical": same,
                          "max_abs_diff": max((a[0].float() - b[0].float()).abs().max().item(),
                                              (a[1].float() - b[1].float()).abs().max().item())}), flush=True)
    for tag, (B, H, P, T, C) in (("config 2 (1024^2)", (1, 24, 4096, 256, 4)), ("config 3 (512^2 x16)", (16, 24, 1024, 256, 2)),
                                 ("config 5 / 8 GPUs", (1, 3, 16384, 256, 32))):
        N = P + T + C
        q, k, v = (torch.randn(B, H, N, 128, device="cuda").to(BF16) for _ in range(3))
        t0, t14 = timed(0, q, k, v, P, T, C), timed(14, q, k, v, P, T, C)
        fl = 4.0 * B * H * (P + T) * (P + T) * 128
        print(json.dumps({"shape": tag, "default_ms": round(t0, 4), "persistent_ms": round(t14, 4),
                          "default_TFLOPs": round(fl / t0 / 1e9, 1), "persistent_TFLOPs": round(fl / t14 / 1e9, 1)}), flush=True)
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
