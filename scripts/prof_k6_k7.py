#!/usr/bin/env python
"""Two launches of each K7 kernel, the grouped K1b and K6 at config-5 (2-GPU) shapes, for one
`ncu --set full` capture (profiles/r01_k6_k7_raw.csv)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from conceptattention_b200 import _cabi, ops  # noqa: E402

BF16 = torch.bfloat16
dev = "cuda:0"
W, Hl, Pl, Tl, C = 2, 12, 8192, 128, 32
H, Nl = W * Hl, Pl + Tl + C
recv = torch.randn(W, 3, Hl, Nl, 128, device=dev).to(BF16)
o_img = torch.randn(1, W * Pl, Hl * 128, device=dev).to(BF16)
o_tc = torch.randn(1, W * Tl + C, Hl * 128, device=dev).to(BF16)
back = torch.randn(W, Nl, Hl * 128, device=dev).to(BF16)
a = torch.randn(1, Pl, 3 * H * 128, device=dev).to(BF16)
b = torch.randn(1, Tl + C, 3 * H * 128, device=dev).to(BF16)
sc = [(torch.rand(128, device=dev) + 0.5).to(BF16) for _ in range(4)]
rope = torch.randn(Nl, 64, 2, device=dev)
n, nclass = 64_000_000, 21
tg = torch.randint(-1, nclass, (n,), device=dev).float()
pr = torch.randint(0, nclass, (n,), device=dev).float()
counts = torch.empty(2 + 3 * nclass, device=dev, dtype=torch.int64)
lib = _cabi.load()
torch.cuda.synchronize()
for _ in range(2):
    ops.seq_unpack_qkv(recv, 1, Pl, Tl, C)
    ops.seq_pack_out(o_img, o_tc, W, Pl, Tl, C)
    ops.seq_unpack_out(back, Pl, Tl, C)
    ops.qkv_prep_grouped(a, b, (sc[0], sc[1]), (sc[2], sc[3]), rope, H, Hl)
    lib.ca_segmentation_counts(pr.data_ptr(), tg.data_ptr(), n, nclass, counts.data_ptr(),
                               torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
print("ok")
