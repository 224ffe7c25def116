#!/usr/bin/env python
"""Micro-benchmark of the K1 kernels (LN+modulate, gated residual+LN, QKV prep) and K5
(finalize / argmax / binary masks) against the HBM roofline.

Inputs rotate over enough distinct buffers (> 2x the 126 MB L2) that every launch streams
from HBM; time = CUDA events over `iters` launches; bytes = the ALGORITHMIC traffic stated in
DESIGN.md section 3 (each tensor read or written once); frac = GB/s / measured copy bandwidth."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from conceptattention_b200 import _cabi, ops  # noqa: E402

BF16 = torch.bfloat16


def timeit(fn, nbuf, iters=40):
    for i in range(nbuf):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        fn(i % nbuf)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e-3


def main():
    peak = 6650.0
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        peak = json.load(open(path))["hbm_gbs"]
    dev = "cuda:0"
    D, H = 3072, 24
    rows_out = []

    def report(name, shape, nbytes, sec):
        gbs = nbytes / sec / 1e9
        rows_out.append((name, shape, sec * 1e6, gbs, gbs / peak))
        print(json.dumps({"kernel": name, "shape": shape, "us": round(sec * 1e6, 2), "GB/s": round(gbs, 1),
                          "frac_hbm": round(gbs / peak, 3)}), flush=True)

    for tag, N in (("cfg2 img stream (N=4096)", 4096), ("cfg2 single stream (N=4352)", 4352),
                   ("cfg5 img stream (N=16384)", 16384)):
        nbuf = max(3, int(400e6 // (N * D * 2)) + 1)
        xs = [torch.randn(1, N, D, device=dev).to(BF16) for _ in range(nbuf)]
        ys = [torch.randn(1, N, D, device=dev).to(BF16) for _ in range(nbuf)]
        outs = [torch.empty(1, N, D, device=dev, dtype=BF16) for _ in range(nbuf)]
        mod = (torch.randn(1, 6 * D, device=dev) * 0.1).to(BF16)
        report("K1a ca_ln_modulate", tag, 2 * N * D * 2,
               timeit(lambda i: ops.ln_modulate(xs[i], mod, None, None, 0, D, out=outs[i]), nbuf))
        report("K1c ca_gate_residual_ln_modulate (x+=g*y, LN out)", tag, 4 * N * D * 2,
               timeit(lambda i: ops.gate_residual_ln_modulate(xs[i], ys[i], mod, None, None, 2 * D, 3 * D, 4 * D,
                                                              out=outs[i]), nbuf))
        report("K1c ca_gate_residual_ln_modulate (x+=g*y only)", tag, 3 * N * D * 2,
               timeit(lambda i: ops.gate_residual_ln_modulate(xs[i], ys[i], mod, None, None, 5 * D, 0, 0,
                                                              want_out=False), nbuf))
        del xs, ys, outs
        nbq = max(3, int(400e6 // (N * 3 * D * 2)) + 1)
        qkvs = [torch.randn(1, N, 3 * D, device=dev).to(BF16) for _ in range(nbq)]
        sc = (torch.ones(128, device=dev, dtype=BF16), torch.ones(128, device=dev, dtype=BF16))
        rope = torch.randn(N, 64, 2, device=dev)
        report("K1b ca_qkv_prep", tag, 2 * N * 3 * D * 2 + N * 512,
               timeit(lambda i: ops.qkv_prep(qkvs[i], None, sc, None, rope, H), nbq))
        report("K1b ca_qkv_prep (+q_pre)", tag, 2 * N * 3 * D * 2 + N * D * 2 + N * 512,
               timeit(lambda i: ops.qkv_prep(qkvs[i], None, sc, None, rope, H, want_q_pre=True), nbq))
        del qkvs
        torch.cuda.empty_cache()
    # K5 at the encode sweep's batch (config 3: 256 images x 2 concepts x 32x32) and at 2048^2 x 32 concepts
    for tag, B, C, side in (("cfg3 256 x C=2 x 32^2", 256, 2, 32), ("cfg5 C=32 x 128^2 x 64 maps", 64, 32, 128)):
        P = side * side
        nbuf = max(3, int(300e6 // (B * C * P * 4)) + 1)
        accs = [torch.rand(B, C, P, device=dev) for _ in range(nbuf)]
        heats = [a.view(B, C, side, side) for a in accs]
        report("K5 ca_finalize_heatmaps", tag, 2 * B * C * P * 4, timeit(lambda i: ops.finalize_heatmaps(accs[i], 16), nbuf))
        report("K5 ca_argmax_mask", tag, B * C * P * 4 + B * P * 8, timeit(lambda i: ops.argmax_mask(heats[i]), nbuf))
        report("K5 ca_binary_masks", tag, B * C * P * 5, timeit(lambda i: ops.binary_masks(heats[i]), nbuf))
        report("K5 ca_gaussian_blur3", tag, 2 * B * C * P * 4, timeit(lambda i: ops.gaussian_blur3(heats[i]), nbuf))
    print("\n| kernel | shape | us | GB/s | frac of measured HBM |\n|---|---|---|---|---|")
    for n, s, us, g, f in rows_out:
        print(f"| {n} | {s} | {us:.1f} | {g:.0f} | {f:.3f} |")


if __name__ == "__main__":
    main()
