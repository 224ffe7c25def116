#!/usr/bin/env python
"""Run one K2 impl on one shape a few times in a fresh process (a launch failure kills the context).
usage: k2_probe.py IMPL B H P T C [iters]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from conceptattention_b200 import ops  # noqa: E402

impl, B, H, P, T, C = (int(a) for a in sys.argv[1:7])
iters = int(sys.argv[7]) if len(sys.argv) > 7 else 20
N = P + T + C
q, k, v = (torch.randn(B, H, N, 128, device="cuda").to(torch.bfloat16) for _ in range(3))
ri, rt = ops.joint_attention(q, k, v, P, T, C, 0, impl=1)
for _ in range(iters):
    oi, ot = ops.joint_attention(q, k, v, P, T, C, 0, impl=impl)
torch.cuda.synchronize()
print("ok", sys.argv[1:7], float((oi.float() - ri.float()).abs().max()), float((ot.float() - rt.float()).abs().max()))
