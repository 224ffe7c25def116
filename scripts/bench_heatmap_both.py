#!/usr/bin/env python
"""One launch for both heatmap spaces of a layer (ca_heatmap_accumulate_both) against the two
launches it replaces and a device copy of the same bytes.  Rotating inputs larger than L2."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from conceptattention_b200 import ops  # noqa: E402

SHAPES = [("cfg2 1024^2 C=4", 1, 4096, 256, 4), ("cfg4 1024^2 C=20", 1, 4096, 512, 20),
          ("cfg3 512^2 C=2 batch 16", 16, 1024, 256, 2), ("cfg5 2048^2 C=32", 1, 16384, 256, 32)]


def timed(fn, n):
    for i in range(3):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n


def main():
    peak = 6458.4
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        peak = json.load(open(path))["hbm_gbs"]
    H, D = 24, 3072
    for name, B, P, T, C in SHAPES:
        nbuf = max(2, int(300e6 // (B * P * D * 4)) + 1)
        sets = []
        for _ in range(nbuf):
            q = torch.randn(B, H, P + T + C, 128, device="cuda").to(torch.bfloat16) * 0.2
            sets.append((torch.randn(B, P, D, device="cuda").to(torch.bfloat16) * 0.2,
                         torch.randn(B, C, D, device="cuda").to(torch.bfloat16) * 0.2, q[:, :, :P], q[:, :, P + T:]))
        acc = torch.zeros(2, B, C, P, device="cuda")
        nbytes = 2 * B * (2 * P * D + 2 * C * D + 8 * C * P)

        def two(i):
            a, b, c, d = sets[i % nbuf]
            ops.heatmap_accumulate(a, b, acc[0], "softmax")
            ops.heatmap_accumulate(c, d, acc[1], "softmax", per_head=True)

        def both(i):
            a, b, c, d = sets[i % nbuf]
            ops.heatmap_accumulate_both(a, b, acc[0], c, d, acc[1], "softmax")
        src = torch.empty(nbytes // 2, dtype=torch.uint8, device="cuda")
        dsts = [torch.empty_like(src) for _ in range(nbuf)]
        t2, t1 = timed(two, 10 * nbuf), timed(both, 10 * nbuf)
        tc = timed(lambda i: dsts[i % nbuf].copy_(src), 10 * nbuf)
        print(json.dumps({"shape": name, "algorithmic_bytes_both_spaces": nbytes, "two_launches_us": round(t2, 2),
                          "one_launch_us": round(t1, 2), "same_bytes_copy_us": round(tc, 2),
                          "one_launch_GBps": round(nbytes / t1 / 1e3, 1), "frac": round(nbytes / t1 / 1e3 / peak, 3)}),
              flush=True)
        del sets, src, dsts
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
