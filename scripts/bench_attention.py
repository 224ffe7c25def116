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


def bench_sdpa(name, B, H, P, T, C, iters, pk):
    """The op K2 replaces, as the reference runs it on this GPU: two
    F.scaled_dot_product_attention calls per double block (modified_double_stream_block.py:112-116
    over [txt;img] and :162-166 over [concept;img], whose image rows are then discarded), one per
    single block (flux/src/flux/math.py:6-12).  bf16, no mask, torch's default backend choice plus
    each backend forced.  `TFLOPs_executed` counts what the library kernel computes,
    `TFLOPs_algorithmic` credits only K2's FLOP count (what the path needs) over the time of BOTH calls."""
    import torch.nn.functional as F
    from torch.nn.attention import SDPBackend, sdpa_kernel
    N1, N2 = T + P, C + P
    mk = lambda n: [[torch.randn(B, H, n, 128, device="cuda").to(torch.bfloat16) for _ in range(3)] for _ in range(2)]
    a1, a2 = mk(N1), (mk(N2) if C > 0 else None)
    need = 4 * 128 * H * B * (N1 ** 2 + C * (C + P))

    def timed(fn):
        for i in range(3):
            fn(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(iters):
            fn(i)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters

    backends = [("default", None), ("flash", SDPBackend.FLASH_ATTENTION), ("cudnn", SDPBackend.CUDNN_ATTENTION),
                ("efficient", SDPBackend.EFFICIENT_ATTENTION)]
    for bname, be in backends:
        try:
            ctx = (lambda: sdpa_kernel([be])) if be is not None else None
            def run1(i):
                return F.scaled_dot_product_attention(*a1[i % 2])
            def run2(i):
                return F.scaled_dot_product_attention(*a2[i % 2])
            if ctx is not None:
                with ctx():
                    ms1 = timed(run1)
                    ms2 = timed(run2) if C > 0 else 0.0
            else:
                ms1 = timed(run1)
                ms2 = timed(run2) if C > 0 else 0.0
        except Exception as e:                                  # backend not available for this shape
            print(json.dumps({"kernel": "F.scaled_dot_product_attention", "backend": bname, "shape": name,
                              "error": f"{type(e).__name__}: {e}"[:160]}), flush=True)
            continue
        tf1 = 4 * 128 * H * B * N1 ** 2 / ms1 / 1e9
        tf2 = 4 * 128 * H * B * N2 ** 2 / ms2 / 1e9 if C > 0 else None
        alg = need / (ms1 + ms2) / 1e9
        print(json.dumps({"kernel": "F.scaled_dot_product_attention", "backend": bname, "shape": name, "B": B,
                          "H": H, "P": P, "T": T, "C": C, "ms_txt_img": round(ms1, 4),
                          "ms_concept_img": round(ms2, 4), "ms_both": round(ms1 + ms2, 4),
                          "TFLOPs_executed_txt_img": round(tf1, 1),
                          "TFLOPs_executed_concept_img": None if tf2 is None else round(tf2, 1),
                          "TFLOPs_algorithmic": round(alg, 1),
                          "frac_sustained_peak_txt_img": round(tf1 / pk["bf16_tflops_sustained"], 3),
                          "frac_sustained_peak_algorithmic": round(alg / pk["bf16_tflops_sustained"], 3)}),
              flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sdpa", action="store_true", help="also time the reference's SDPA calls on the same shapes")
    ap.add_argument("--impls", default="0,1,2,3", help="CA_K2_* per-call kernel ids: 0 auto, 1 v3, 2 v7, 3 v11")
    ap.add_argument("--only", default="")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--interleave", default="", choices=["", "gemm", "copy"],
                    help="launch another kernel before every K2 launch (what a forward does) and time K2 alone "
                         "with events around each launch")
    args = ap.parse_args()
    pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(
        os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}
    lib = _cabi.load()
    plan_sms = int(os.environ.get("CA_STREAMK", "0"))   # -1: never split the last wave (v3 only)
    for name, B, H, P, T, C in SHAPES:
        if args.only not in name:
            continue
        N = P + T + C
        sets = [[(torch.randn(B, H, N, 128, device="cuda") * s).to(torch.bfloat16) for s in (1.0, 1.0, 1.0)]
                for _ in range(2)]
        flops = 4 * 128 * H * B * ((T + P) ** 2 + C * (C + P))
        for impl in [int(i) for i in args.impls.split(",")]:
            for i in range(3):
                ops.joint_attention(*sets[i % 2], P, T, C, impl=impl, plan_sms=plan_sms)
            torch.cuda.synchronize()
            if args.interleave:
                a = torch.randn(4352, 3072, device="cuda").to(torch.bfloat16)
                w = torch.randn(3072, 9216, device="cuda").to(torch.bfloat16)
                big = torch.empty(160 << 20, dtype=torch.uint8, device="cuda")
                big2 = torch.empty_like(big)
                evs = []
                for i in range(args.iters):
                    if args.interleave == "gemm":
                        torch.mm(a, w)
                    else:
                        big2.copy_(big)
                    ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    ea.record()
                    ops.joint_attention(*sets[i % 2], P, T, C, impl=impl, plan_sms=plan_sms)
                    eb.record()
                    evs.append((ea, eb))
                torch.cuda.synchronize()
                ms = sum(x.elapsed_time(y) for x, y in evs) / args.iters
            else:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for i in range(args.iters):
                    ops.joint_attention(*sets[i % 2], P, T, C, impl=impl, plan_sms=plan_sms)
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / args.iters
            tf = flops / ms / 1e9
            print(json.dumps({"kernel": "ca_joint_attention", "impl": impl, "shape": name, "B": B, "H": H,
                              "P": P, "T": T, "C": C, "ms": round(ms, 4), "TFLOPs": round(tf, 1),
                              "frac_burst_peak": round(tf / pk["bf16_tflops"], 3),
                              "frac_sustained_peak": round(tf / pk["bf16_tflops_sustained"], 3)}), flush=True)
        del sets
        torch.cuda.empty_cache()
        if args.sdpa:
            bench_sdpa(name, B, H, P, T, C, args.iters, pk)
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
