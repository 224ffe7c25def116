#!/usr/bin/env python
"""Is K2's in-pipeline timing data-dependent?  Captures the q, k, v of a few K2 launches of a real
forward (flux-schnell shapes, random-init weights, 1024^2) and times the kernels named by id on exactly
those tensors, alternating, next to randn tensors of the same shape.  usage: k2_capture_ab.py [ids...]"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from conceptattention_b200 import ConceptAttentionFluxPipeline, ops  # noqa: E402
from conceptattention_b200.weights import CONFIGS, init_state_dict  # noqa: E402


def timed(fn, n=60):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def main():
    ids = [int(a) for a in sys.argv[1:]] or [3, 4]
    dev = torch.device("cuda", 0)
    sd = init_state_dict(CONFIGS["flux-schnell"], seed=0, device=dev, dtype=torch.bfloat16)
    pipe = ConceptAttentionFluxPipeline(model_name="flux-schnell", device=dev, state_dict=sd)
    pipe.flux_generator.model.use_graphs = False
    captured, orig, count = [], ops.joint_attention, [0]

    def spy(q, k, v, P, T, C, *a, **kw):
        count[0] += 1
        if count[0] in (2, 10, 19, 30, 57 + 10, 3 * 57 + 19):       # double and single blocks, early and late steps
            captured.append((count[0], q.clone(), k.clone(), v.clone(), P, T, C))
        return orig(q, k, v, P, T, C, *a, **kw)
    ops.joint_attention = spy
    import conceptattention_b200.dit as dit
    dit.ops.joint_attention = spy
    pipe.generate_image("A dragon standing on a rock", ["dragon", "rock", "sky", "sun"], width=1024, height=1024,
                        num_inference_steps=4, return_pil_heatmaps=False, decode_image=False)
    torch.cuda.synchronize()
    ops.joint_attention = dit.ops.joint_attention = orig
    for idx, q, k, v, P, T, C in captured:
        rq, rk, rv = (torch.randn_like(q.float()).to(torch.bfloat16) for _ in range(3))
        stats = {"q_rms": float(q.float().pow(2).mean().sqrt()), "k_rms": float(k.float().pow(2).mean().sqrt()),
                 "score_max": float((q[0, 0, :512].float() @ k[0, 0].float().t()).abs().max() / 128 ** 0.5)}
        row = {"call": idx, "P": P, "T": T, "C": C, **{k_: round(v_, 3) for k_, v_ in stats.items()}}
        for rep in range(2):
            for impl in ids:
                row[f"real impl{impl} run{rep}"] = round(timed(lambda: orig(q, k, v, P, T, C, impl=impl)), 4)
                row[f"randn impl{impl} run{rep}"] = round(timed(lambda: orig(rq, rk, rv, P, T, C, impl=impl)), 4)
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
