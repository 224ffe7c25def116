#!/usr/bin/env python
"""Numerical check of the head-sharded modes over NCCL (the GPU tests use gloo on one GPU):
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 scripts/check_hs_nccl.py
Tiny model, B = 1.  For the sequence-parallel and the tensor-parallel engine: eager call, then the
CUDA-graph replay (third call of the same signature), both against the one-GPU engine on rank 0's GPU.
Prints one JSON line per comparison and exits non-zero on a mismatch."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import golden_inputs as gi  # noqa: E402
from conceptattention_b200.dit import FluxDiTEngine, HeatmapAccumulator  # noqa: E402
from conceptattention_b200.parallel import HeadShard, init_distributed  # noqa: E402
from conceptattention_b200.weights import CONFIGS, init_state_dict  # noqa: E402


def main():
    rank, world, local = init_distributed("nccl")
    dev = f"cuda:{local}"
    torch.set_grad_enabled(False)
    p = CONFIGS["tiny"]
    sd = init_state_dict(p, seed=gi.WEIGHT_SEED, norm_scale_jitter=0.25)
    inp = gi.dit_inputs(p, gi.TINY_SIDE, gi.TINY_T, gi.TINY_C, gi.INPUT_SEED, batch=1)
    inp = {k: v.to(dev) for k, v in inp.items()}
    t = torch.full((1,), 0.75, device=dev)
    P = inp["img"].shape[1]
    buf = torch.zeros(2, 1, gi.TINY_C, P, device=dev)

    def run(engine):
        acc = HeatmapAccumulator(1, gi.TINY_C, P, dev, [0, 2], None, buffer=buf)
        pred, _ = engine(img=inp["img"], img_ids=inp["img_ids"], txt=inp["txt"], txt_ids=inp["txt_ids"],
                         concepts=inp["concepts"], concept_ids=inp["concept_ids"], concept_vec=inp["concept_vec"],
                         y=inp["vec"], timesteps=t, accumulator=acc, collect_vectors=False)
        out, cross = acc.finalize()
        return pred.float().clone(), out.clone(), cross.clone()

    full = run(FluxDiTEngine(p, sd, dev))
    if rank == 0:
        print(json.dumps({"stage": "one-GPU engine done"}), flush=True)
    bad = 0
    for mode, tol in (("sequence", 1e-6), ("tensor", 5e-2)):
        eng = FluxDiTEngine(p, sd, dev, head_shard=HeadShard(rank, world, p.num_heads,
                                                             sequence_parallel=mode == "sequence"))
        eager = run(eng)
        if rank == 0:
            print(json.dumps({"stage": mode + " eager done"}), flush=True)
        run(eng)                                        # second sighting: capture
        if rank == 0:
            print(json.dumps({"stage": mode + " capture done"}), flush=True)
        replay = run(eng)                               # replayed graph
        graphed = len(eng._graphs) > 0
        for tag, got in (("eager", eager), ("replay", replay)):
            rel = [((a - b).abs().max() / b.abs().max()).item() for a, b in zip(got, full)]
            ok = all(r <= tol for r in rel) and (tag == "eager" or graphed)
            bad += not ok
            if rank == 0:
                print(json.dumps({"mode": mode, "path": tag, "graphs": len(eng._graphs), "world": world,
                                  "max_rel_vs_one_gpu": {"pred": rel[0], "output": rel[1], "cross": rel[2]},
                                  "ok": ok}), flush=True)
        same = all(torch.equal(a, b) for a, b in zip(eager, replay))
        if rank == 0:
            print(json.dumps({"mode": mode, "replay_equals_eager": same}), flush=True)
    dist.barrier()
    torch.cuda.synchronize()
    sys.stdout.flush()
    os._exit(1 if bad else 0)     # captured graphs hold NCCL work: skip the orderly teardown


if __name__ == "__main__":
    main()
