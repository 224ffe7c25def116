#!/usr/bin/env python
"""Micro-benchmark of the K7 exchange-layout kernels and the grouped K1b (config 5 on 8 GPUs:
W = 8, 3 heads per rank, 2048 image + 32 text rows per rank, 32 concepts) and of K6
(segmentation counters) against the HBM roofline.  One GPU: the all-to-all itself is not part of
these kernels.  Buffers rotate over more than 2x the L2; bytes = algorithmic (read once + written once)."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from conceptattention_b200 import metrics, ops  # noqa: E402
from scripts.bench_k1 import timeit  # noqa: E402

BF16 = torch.bfloat16


def main():
    peak = 6650.0
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        peak = json.load(open(path))["hbm_gbs"]
    dev = "cuda:0"

    def report(name, shape, nbytes, sec):
        gbs = nbytes / sec / 1e9
        print(json.dumps({"kernel": name, "shape": shape, "us": round(sec * 1e6, 2), "GB/s": round(gbs, 1),
                          "frac_hbm": round(gbs / peak, 3)}), flush=True)

    for W, Hl, Pl, Tl, C in ((8, 3, 2048, 32, 32), (2, 12, 8192, 128, 32)):
        H, Nl, N = W * Hl, Pl + Tl + C, W * (Pl + Tl) + C
        tag = f"W={W} Hl={Hl} Pl={Pl} Tl={Tl} C={C}"
        qkv_bytes = W * 3 * Hl * Nl * 256
        nbuf = max(3, int(300e6 // qkv_bytes) + 1)
        recv = [torch.randn(W, 3, Hl, Nl, 128, device=dev).to(BF16) for _ in range(nbuf)]
        report("K7 ca_seq_unpack_qkv", tag, 2 * qkv_bytes, timeit(lambda i: ops.seq_unpack_qkv(recv[i], 1, Pl, Tl, C), nbuf))
        out_bytes = W * Nl * Hl * 256
        nb2 = max(3, int(300e6 // out_bytes) + 1)
        o_img = [torch.randn(1, W * Pl, Hl * 128, device=dev).to(BF16) for _ in range(nb2)]
        o_tc = [torch.randn(1, W * Tl + C, Hl * 128, device=dev).to(BF16) for _ in range(nb2)]
        report("K7 ca_seq_pack_out", tag, 2 * out_bytes, timeit(lambda i: ops.seq_pack_out(o_img[i], o_tc[i], W, Pl, Tl, C), nb2))
        back = [torch.randn(W, Nl, Hl * 128, device=dev).to(BF16) for _ in range(nb2)]
        report("K7 ca_seq_unpack_out", tag, 2 * out_bytes, timeit(lambda i: ops.seq_unpack_out(back[i], Pl, Tl, C), nb2))
        a = [torch.randn(1, Pl, 3 * H * 128, device=dev).to(BF16) for _ in range(nbuf)]
        b = [torch.randn(1, Tl + C, 3 * H * 128, device=dev).to(BF16) for _ in range(nbuf)]
        sc = [(torch.rand(128, device=dev) + 0.5).to(BF16) for _ in range(4)]
        rope = torch.randn(Nl, 64, 2, device=dev)
        report("K1b ca_qkv_prep_grouped", tag, 2 * Nl * 3 * H * 256,
               timeit(lambda i: ops.qkv_prep_grouped(a[i], b[i], (sc[0], sc[1]), (sc[2], sc[3]), rope, H, Hl), nbuf))
        report("K1b ca_qkv_prep (same rows, plain layout)", tag, 2 * Nl * 3 * H * 256,
               timeit(lambda i: ops.qkv_prep(a[i], b[i], (sc[0], sc[1]), (sc[2], sc[3]), rope, H), nbuf))
        del recv, o_img, o_tc, back, a, b
        torch.cuda.empty_cache()

    for n, nclass in ((224 * 224 * 2 * 64, 2), (9_000_000, 21), (64_000_000, 21), (64_000_000, 150)):
        nbuf = max(3, int(300e6 // (8 * n)) + 1)
        tg = [torch.randint(-1, nclass, (n,), device=dev).float() for _ in range(nbuf)]
        pr = [torch.randint(0, nclass, (n,), device=dev).float() for _ in range(nbuf)]
        counts = torch.empty(2 + 3 * nclass, device=dev, dtype=torch.int64)
        from conceptattention_b200 import _cabi
        lib = _cabi.load()
        st = torch.cuda.current_stream().cuda_stream
        report("K6 ca_segmentation_counts", f"n={n} nclass={nclass}", 8 * n,
               timeit(lambda i: lib.ca_segmentation_counts(pr[i].data_ptr(), tg[i].data_ptr(), n, nclass,
                                                           counts.data_ptr(), st), nbuf))
        del tg, pr
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
