"""Sequence-parallel form of the head-sharded mode: token-wise ops on 1/W of the
rows with replicated weights, heads split only inside the attention (two all-to-alls per block).

* kernel level, one process: the grouped K1b output + K7 permutations, with the all-to-all
  simulated by indexing, must reproduce K1b / K2 on the full row order bit for bit;
* engine level: two processes share cuda:0 over gloo (collectives staged through the host); the
  sequence-parallel engine must reproduce the one-GPU engine (same GEMM rows, same attention
  maths: differences only from cuBLAS choosing another kernel for another M)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import golden_inputs as gi

pytestmark = pytest.mark.gpu
BF16 = torch.bfloat16


@pytest.mark.parametrize("W,Hl,Pl,Tl,C", [(2, 1, 72, 36, 3), (4, 2, 160, 8, 5), (3, 2, 130, 0, 0), (8, 3, 256, 32, 32)])
def test_exchange_kernels_reproduce_full_order(W, Hl, Pl, Tl, C):
    from conceptattention_b200 import ops, _cabi
    H, P, T = W * Hl, W * Pl, W * Tl
    N = P + T + C
    g = torch.Generator().manual_seed(W * 1000 + Pl)
    img_qkv = torch.randn(1, P, 3 * H * 128, generator=g).to(BF16).cuda()
    tc_qkv = torch.randn(1, T + C, 3 * H * 128, generator=g).to(BF16).cuda() if T + C else None
    sc = [(torch.rand(128, generator=g) + 0.5).to(BF16).cuda() for _ in range(4)]
    rope = torch.randn(N, 64, 2, generator=g).cuda()
    q, k, v, q_pre = ops.qkv_prep(img_qkv, tc_qkv, (sc[0], sc[1]), (sc[2], sc[3]) if tc_qkv is not None else None,
                                  rope, H, want_q_pre=True)
    # plan_sms=-1: whole units per CTA, so a head slice reproduces the all-heads launch bit for bit
    # (with key-step ranges the fp32 merge of a cut unit differs from in-TMEM accumulation by a bf16 ulp)
    ref_img, ref_tc = ops.joint_attention(q, k, v, P, T, C, plan_sms=-1)

    sends, qpres = [], []
    for r in range(W):                                   # what rank r computes from its rows
        a = img_qkv[:, r * Pl:(r + 1) * Pl]
        b = None
        if tc_qkv is not None:
            b = torch.cat((tc_qkv[:, r * Tl:(r + 1) * Tl], tc_qkv[:, T:]), 1).contiguous()
        rl = torch.cat((rope[r * Pl:(r + 1) * Pl], rope[P + r * Tl:P + (r + 1) * Tl], rope[P + T:]), 0).contiguous()
        send, qp = ops.qkv_prep_grouped(a.contiguous(), b, (sc[0], sc[1]), (sc[2], sc[3]) if b is not None else None,
                                        rl, H, Hl, want_q_pre=True)
        sends.append(send)
        qpres.append(qp)
    outs = []
    for r in range(W):                                   # all-to-all: rank r receives chunk r of every rank
        recv = torch.stack([sends[s][r] for s in range(W)]).contiguous()
        ql, kl, vl = ops.seq_unpack_qkv(recv, r, Pl, Tl, C)
        for got, full in ((ql, q), (kl, k), (vl, v)):
            assert torch.equal(got, full[:, r * Hl:(r + 1) * Hl]), r
        o_img, o_tc = ops.joint_attention(ql, kl, vl, P, T, C, plan_sms=-1)
        assert torch.equal(o_img, ref_img[:, :, r * Hl * 128:(r + 1) * Hl * 128])
        outs.append(ops.seq_pack_out(o_img, o_tc, W, Pl, Tl, C))
    for r in range(W):
        back = torch.stack([outs[s][r] for s in range(W)]).contiguous()
        a_img, a_tc = ops.seq_unpack_out(back, Pl, Tl, C)
        assert torch.equal(a_img, ref_img[:, r * Pl:(r + 1) * Pl])
        want_tc = torch.cat((ref_tc[:, r * Tl:(r + 1) * Tl], ref_tc[:, T:]), 1)
        assert torch.equal(a_tc, want_tc)
        # caller-owned output rows (the single-block form)
        buf = torch.empty(1, Pl + Tl + C, H * 128, device="cuda", dtype=BF16)
        ops.seq_unpack_out(back, Pl, Tl, C, out=(buf[:, :Pl], buf[:, Pl:]))
        assert torch.equal(buf, torch.cat((a_img, a_tc), 1))
        # pre-RoPE q of the local rows, all heads
        want_qp = torch.cat((q_pre[:, :, r * Pl:(r + 1) * Pl], q_pre[:, :, P + r * Tl:P + (r + 1) * Tl],
                             q_pre[:, :, P + T:]), 2)
        assert torch.equal(qpres[r], want_qp)
    lib = _cabi.load()
    assert lib.ca_seq_unpack_qkv(0, 2, 0, 1, 8, 0, 0, 0, 0, 0, 0) != 0          # null pointers refused
    assert lib.ca_seq_unpack_qkv(1 << 8, 2, 5, 1, 8, 0, 0, 1 << 8, 1 << 8, 1 << 8, 0) != 0   # self outside world


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from conceptattention_b200.dit import FluxDiTEngine, HeatmapAccumulator
        from conceptattention_b200.parallel import HeadShard
        from conceptattention_b200.weights import CONFIGS, init_state_dict
        torch.set_grad_enabled(False)
        p = CONFIGS["tiny"]
        sd = init_state_dict(p, seed=gi.WEIGHT_SEED, norm_scale_jitter=0.25)
        inp = gi.dit_inputs(p, gi.TINY_SIDE, gi.TINY_T, gi.TINY_C, gi.INPUT_SEED, batch=1)
        t = torch.full((1,), 0.75)
        P = inp["img"].shape[1]
        kw = dict(img=inp["img"], img_ids=inp["img_ids"], txt=inp["txt"], txt_ids=inp["txt_ids"],
                  concepts=inp["concepts"], concept_ids=inp["concept_ids"], concept_vec=inp["concept_vec"],
                  y=inp["vec"], timesteps=t)

        def run(engine, **extra):
            acc = HeatmapAccumulator(1, gi.TINY_C, P, "cuda:0", [0, 2], None)
            pred, _ = engine(**kw, accumulator=acc, collect_vectors=False, **extra)
            out, cross = acc.finalize()
            return (None if pred is None else pred.float().cpu()), out.cpu(), cross.cpu()

        eng = FluxDiTEngine(p, sd, "cuda:0", head_shard=HeadShard(rank, world, p.num_heads, sequence_parallel=True))
        sp = run(eng)
        sp_stop = run(eng, stop_after_multimodal_attentions=True)        # encode path (dead tail skipped)
        _, vec_sp = eng(**kw, collect_vectors=True)
        with pytest.raises(ValueError):
            eng(**{**kw, "img": inp["img"][:, :P - 1], "img_ids": inp["img_ids"][:, :P - 1]}, collect_vectors=False)
        if rank == 0:
            full = FluxDiTEngine(p, sd, "cuda:0")
            ret["sp"], ret["sp_stop"] = sp, sp_stop
            ret["full"] = run(full)
            _, vec_full = full(**kw, collect_vectors=True)
            ret["vec_sp"] = {k: v.float().cpu() for k, v in vec_sp.items()}
            ret["vec_full"] = {k: v.float().cpu() for k, v in vec_full.items()}
        else:
            ret["sp1"] = sp
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_sequence_parallel_engine_matches_one_gpu():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(2, port, ret), nprocs=2, join=True)
    for a, b in zip(ret["sp"], ret["sp1"]):                     # every rank ends with the same result
        torch.testing.assert_close(a, b, rtol=0, atol=0)
    mae = lambda a, b: (a - b).abs().mean().item()
    pred_s, out_s, cross_s = ret["sp"]
    pred_f, out_f, cross_f = ret["full"]
    print("pred max rel", ((pred_s - pred_f).abs().max() / pred_f.abs().max()).item(),
          "heatmap mae", mae(out_s, out_f), mae(cross_s, cross_f))
    assert ((pred_s - pred_f).abs().max() / pred_f.abs().max()).item() < 2e-2
    for got, ref in ((out_s, out_f), (cross_s, cross_f)):
        assert (got.argmax(1) == ref.argmax(1)).float().mean().item() >= 0.999
        assert mae(got, ref) < 1e-3
    for got, ref in zip(ret["sp_stop"][1:], (out_f, cross_f)):  # heatmaps of the encode path
        assert mae(got, ref) < 1e-3
    assert ret["sp_stop"][0] is None
    for k, ref in ret["vec_full"].items():                      # gathered raw vectors, reference layout
        got = ret["vec_sp"][k]
        assert got.shape == ref.shape, k
        assert ((got - ref).abs().max() / ref.abs().max()).item() < 3e-2, k
