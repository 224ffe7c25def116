"""Head-sharded modes over NCCL on two real GPUs (one process per GPU) — the hardware version of
test_gpu_headshard.py / test_gpu_seqpar.py, which emulate two ranks on one GPU over gloo.
Skipped when fewer than 2 GPUs are visible (`gpurun --gpus 2 -- python -m pytest tests/test_gpu_nccl.py`).

Sequence-parallel form: the same GEMM rows and the same per-head attention are computed, only on
another rank -> prediction bit-identical to one GPU, heatmaps to fp32 round-off.  Tensor-parallel
form: row-parallel partial sums are a different bf16 evaluation -> held to the bf16 noise floor
(no further from the fp32 oracle than 1.5x the one-GPU engine).  Both: CUDA-graph replay (NCCL
collectives captured) bit-identical to the eager run."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import golden_inputs as gi

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    dev = f"cuda:{rank}"
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device(dev))
    from conceptattention_b200.dit import FluxDiTEngine, HeatmapAccumulator
    from conceptattention_b200.parallel import HeadShard
    from conceptattention_b200.weights import CONFIGS, init_state_dict
    torch.set_grad_enabled(False)
    p = CONFIGS["tiny"]
    sd = init_state_dict(p, seed=gi.WEIGHT_SEED, norm_scale_jitter=0.25)
    inp = {k: v.to(dev) for k, v in gi.dit_inputs(p, gi.TINY_SIDE, gi.TINY_T, gi.TINY_C, gi.INPUT_SEED).items()}
    t = torch.full((1,), 0.75, device=dev)
    P = inp["img"].shape[1]
    buf = torch.zeros(2, 1, gi.TINY_C, P, device=dev)

    def run(engine):
        acc = HeatmapAccumulator(1, gi.TINY_C, P, dev, [0, 2], None, buffer=buf)
        pred, _ = engine(img=inp["img"], img_ids=inp["img_ids"], txt=inp["txt"], txt_ids=inp["txt_ids"],
                         concepts=inp["concepts"], concept_ids=inp["concept_ids"],
                         concept_vec=inp["concept_vec"], y=inp["vec"], timesteps=t, accumulator=acc,
                         collect_vectors=False)
        out, cross = acc.finalize()
        return pred.float().cpu(), out.cpu().clone(), cross.cpu().clone()

    res = {"full": run(FluxDiTEngine(p, sd, dev))}
    for mode in ("sequence", "sequence_nccl", "tensor"):
        # "sequence": exchanges through peer memory (symmetric allocation + barrier + gather kernels)
        # when torch can set it up, "sequence_nccl": the NCCL all-to-all form of the same exchange
        eng = FluxDiTEngine(p, sd, dev, head_shard=HeadShard(
            rank, world, p.num_heads, sequence_parallel=mode.startswith("sequence"),
            peer_exchange=False if mode == "sequence_nccl" else None))
        res[mode + "_eager"] = run(eng)
        res[mode + "_peer"] = eng.hs._symm is not None
        run(eng)                                           # second sighting: capture
        res[mode + "_replay"] = run(eng)
        res[mode + "_graphs"] = len(eng._graphs)
    if rank == 0:
        import oracle
        f32 = lambda d: {k: v.cpu().to(torch.bfloat16).float() for k, v in d.items()}
        ii, sdd = f32(inp), f32(sd)
        _, vecs = oracle.dit_forward(sdd, p.asdict(), ii["img"], ii["img_ids"], ii["txt"], ii["txt_ids"],
                                     ii["concepts"], ii["concept_ids"], ii["concept_vec"],
                                     t.cpu().to(torch.bfloat16).float(), ii["vec"], None)
        res["oracle"] = tuple(
            oracle.heatmaps_from_vectors(vecs[a][None], vecs[b][None], [0, 2], [0])
            for a, b in (("output_space_image_vectors", "output_space_concept_vectors"),
                         ("cross_attention_image_vectors", "cross_attention_concept_vectors")))
        ret.update(res)
    dist.barrier()
    torch.cuda.synchronize()
    os._exit(0)            # captured graphs hold NCCL work: skip the orderly teardown


def test_head_sharded_modes_over_nccl_on_two_gpus():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(2, port, ret), nprocs=2, join=True)
    full, ora = ret["full"], ret["oracle"]
    rel = lambda a, b: ((a - b).abs().max() / b.abs().max()).item()
    mae = lambda a, b: (a - b).abs().mean().item()
    print("peer-memory exchange active:", ret["sequence_peer"])
    assert not ret["sequence_nccl_peer"]
    for mode in ("sequence", "sequence_nccl", "tensor"):
        eager, replay = ret[mode + "_eager"], ret[mode + "_replay"]
        assert ret[mode + "_graphs"] >= 1, mode
        for a, b in zip(eager, replay):
            assert torch.equal(a, b), mode                # graph replay == eager, bit for bit
        d = [rel(a, b) for a, b in zip(eager, full)]
        print(f"NCCL 2 GPUs [{mode}-parallel] vs one-GPU engine: pred {d[0]:.3e} output {d[1]:.3e} cross {d[2]:.3e}")
        if mode.startswith("sequence"):
            assert torch.equal(eager[0], full[0])          # same rows, same GEMMs, elsewhere
            assert d[1] < 1e-5 and d[2] < 1e-5
        else:
            for got, ref, o, tag in ((eager[1], full[1], ora[0], "output"), (eager[2], full[2], ora[1], "cross")):
                assert mae(got, o) <= 1.5 * mae(ref, o) + 1e-4, (tag, mae(got, o), mae(ref, o))
                assert (got.argmax(1) == ref.argmax(1)).float().mean().item() >= 0.999
            assert d[0] < 5e-2
