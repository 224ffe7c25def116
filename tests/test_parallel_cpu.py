"""Host-side multi-GPU logic on CPU: world_size-2 gloo process groups (no GPU needed).
Covers the data-parallel work split + result gather and the head-shard collectives /
weight slicing of conceptattention_b200.parallel."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
from conceptattention_b200.parallel import HeadShard, gather_results, shard_indices


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, fn, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ret[rank] = fn(rank, world)
    finally:
        dist.destroy_process_group()


def _run(fn, world=2):
    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(world, _free_port(), fn, ret), nprocs=world, join=True)
    return dict(ret)


def test_shard_indices_partition():
    for n in (0, 1, 7, 256):
        for world in (1, 2, 4, 8):
            parts = [shard_indices(n, r, world) for r in range(world)]
            assert sorted(sum(parts, [])) == list(range(n))
            assert max(map(len, parts)) - min(map(len, parts)) <= 1


def _dp_job(rank, world):
    n = 7                                                  # ragged: ranks get 4 and 3 items
    mine = shard_indices(n, rank, world)
    local = torch.stack([torch.full((2, 3), float(i)) for i in mine]) if mine else torch.zeros(0, 2, 3)
    out = gather_results(local, n, rank, world, dst=0)
    return None if out is None else out[:, 0, 0].tolist()


def test_data_parallel_gather_restores_item_order():
    ret = _run(_dp_job)
    assert ret[0] == [float(i) for i in range(7)] and ret[1] is None


def _head_job(rank, world):
    H, d, N, C = 4, 128, 6, 3
    hs = HeadShard(rank, world, H)
    g = torch.Generator().manual_seed(0)
    full = torch.randn(2, N, H * d, generator=g)                       # "B L (H D)"
    mine = full[..., hs.h0 * d:hs.h1 * d]
    ok_gather = torch.equal(hs.all_gather_heads(mine.clone()), full)
    qpre = torch.randn(2, H, N, d, generator=g)                        # [B,H,N,d]
    ok_gather1 = torch.equal(hs.all_gather(qpre[:, hs.h0:hs.h1].clone(), 1), qpre)
    # partial logits over the local head slice sum to the full logits (linearity over H*d)
    img, con = full, torch.randn(2, C, H * d, generator=g)
    part = torch.einsum("bpk,bck->bcp", mine, con[..., hs.h0 * d:hs.h1 * d])
    tot = hs.all_reduce_sum(part.clone())
    ref = torch.einsum("bpk,bck->bcp", img, con)
    ok_sum = torch.allclose(tot, ref, rtol=1e-5, atol=1e-4)
    # qkv row slicing: sliced fused weight == the rank's heads of the full projection
    w = torch.randn(3 * H * d, 16, generator=g)
    x = torch.randn(5, 16, generator=g)
    q, k, v = oracle.split_heads((x @ w.t())[None], H)
    ql, kl, vl = oracle.split_heads((x @ w[hs.qkv_rows(d)].t())[None], hs.h_local)
    ok_rows = all(torch.equal(a[:, hs.h0:hs.h1], b) for a, b in ((q, ql), (k, kl), (v, vl)))
    return ok_gather, ok_gather1, ok_sum, ok_rows


def test_head_shard_collectives_and_weight_slices():
    ret = _run(_head_job)
    for r in (0, 1):
        assert ret[r] == (True, True, True, True), (r, ret[r])


def _tp_job(rank, world):
    """Sharded linears reproduce the full ones: column-parallel then row-parallel MLP, the
    head-sliced proj, and the single block's linear2 over cat(attn, gelu(mlp))."""
    import torch.nn.functional as F
    from conceptattention_b200.weights import CONFIGS, init_state_dict
    p = CONFIGS["tiny"]
    D, M, d = p.hidden_size, int(p.hidden_size * p.mlp_ratio), p.hidden_size // p.num_heads
    full = init_state_dict(p, seed=3)
    for i in range(p.depth_single_blocks):                # the engine's split of linear1 / linear2
        w = full.pop(f"single_blocks.{i}.linear2.weight")
        full[f"single_blocks.{i}.linear2a.weight"], full[f"single_blocks.{i}.linear2b.weight"] = w[:, :D], w[:, D:]
        w1, b1 = full.pop(f"single_blocks.{i}.linear1.weight"), full.pop(f"single_blocks.{i}.linear1.bias")
        full[f"single_blocks.{i}.linear1q.weight"], full[f"single_blocks.{i}.linear1q.bias"] = w1[:3 * D], b1[:3 * D]
        full[f"single_blocks.{i}.linear1m.weight"], full[f"single_blocks.{i}.linear1m.bias"] = w1[3 * D:], b1[3 * D:]
    hs = HeadShard(rank, world, p.num_heads)
    mine = hs.shard_state_dict(dict(full), p)
    g = torch.Generator().manual_seed(1)
    x, attn = torch.randn(5, D, generator=g), torch.randn(5, D, generator=g)
    lin = lambda sd, k, t: F.linear(t, sd[k + ".weight"], sd.get(k + ".bias"))
    pre = "double_blocks.1.txt"
    ref_mlp = lin(full, pre + "_mlp.2", F.gelu(lin(full, pre + "_mlp.0", x), approximate="tanh"))
    got_mlp = hs.all_reduce_sum(lin(mine, pre + "_mlp.2", F.gelu(lin(mine, pre + "_mlp.0", x), approximate="tanh")))
    ref_proj = lin(full, pre + "_attn.proj", attn)
    got_proj = hs.all_reduce_sum(lin(mine, pre + "_attn.proj", attn[:, hs.h0 * d:hs.h1 * d]))
    sp = "single_blocks.0"
    gl = F.gelu(lin(full, sp + ".linear1m", x), approximate="tanh")
    ref_l2 = F.linear(torch.cat((attn, gl), 1), torch.cat((full[sp + ".linear2a.weight"], full[sp + ".linear2b.weight"]), 1),
                      full[sp + ".linear2.bias"])
    part = F.linear(attn[:, hs.h0 * d:hs.h1 * d], mine[sp + ".linear2a.weight"], mine[sp + ".linear2.bias"])
    part = part + F.gelu(lin(mine, sp + ".linear1m", x), approximate="tanh") @ mine[sp + ".linear2b.weight"].t()
    got_l2 = hs.all_reduce_sum(part)
    ok_bf16 = torch.allclose(hs.all_reduce_sum(torch.full((4,), 1.5, dtype=torch.bfloat16)),
                             torch.full((4,), 1.5 * world, dtype=torch.bfloat16))
    return tuple(torch.allclose(a, b, rtol=1e-4, atol=1e-5) for a, b in
                 ((got_mlp, ref_mlp), (got_proj, ref_proj), (got_l2, ref_l2))) + (ok_bf16,)


def test_tensor_parallel_linears_match_full():
    ret = _run(_tp_job)
    for r in (0, 1):
        assert ret[r] == (True, True, True, True), (r, ret[r])


def test_head_shard_rejects_uneven_split():
    import pytest
    with pytest.raises(ValueError):
        HeadShard(0, 5, 24)


def _seq_job(rank, world):
    """Sequence-parallel exchange on CPU: local rows x all heads -> all-to-all -> local heads x all
    rows -> attention -> all-to-all back -> local rows x all heads (K7's permutations in plain torch)."""
    import torch.nn.functional as F
    H, d, Pl, Tl = 4, 16, 6, 2
    W, Hl, Nl = world, H // world, Pl + Tl
    hs = HeadShard(rank, world, H, sequence_parallel=True)
    g = torch.Generator().manual_seed(1)
    # full problem in K2 row order [img (W*Pl) | txt (W*Tl)], bf16 like the device path
    qkv = torch.randn(3, H, W * Nl, d, generator=g).to(torch.bfloat16)
    rows = torch.cat((torch.arange(rank * Pl, (rank + 1) * Pl), W * Pl + torch.arange(rank * Tl, (rank + 1) * Tl)))
    local = qkv[:, :, rows]                                            # [3,H,Nl,d] what this rank computed
    send = local.view(3, W, Hl, Nl, d).permute(1, 0, 2, 3, 4).contiguous()   # [W,3,Hl,Nl,d]
    recv = hs.all_to_all(send)
    recv2, wait = hs.all_to_all_begin(send)
    wait()
    ok_begin = torch.equal(recv, recv2)
    full_rows = torch.cat([torch.cat((torch.arange(w * Pl, (w + 1) * Pl), W * Pl + torch.arange(w * Tl, (w + 1) * Tl)))
                           for w in range(W)])
    mine = torch.empty(3, Hl, W * Nl, d, dtype=torch.bfloat16)
    mine[:, :, full_rows] = recv.permute(1, 2, 0, 3, 4).reshape(3, Hl, W * Nl, d)
    ok_qkv = torch.equal(mine, qkv[:, rank * Hl:(rank + 1) * Hl])
    out = F.scaled_dot_product_attention(mine[0].float(), mine[1].float(), mine[2].float())      # [Hl,N,d]
    back = out[:, full_rows].view(Hl, W, Nl, d).permute(1, 2, 0, 3).contiguous()                 # [W,Nl,Hl,d]
    got = hs.all_to_all(back).permute(1, 0, 2, 3).reshape(Nl, H * d)                             # rows x all heads
    ref = F.scaled_dot_product_attention(qkv[0].float(), qkv[1].float(), qkv[2].float())         # [H,N,d]
    ok_out = torch.equal(got, ref[:, rows].permute(1, 0, 2).reshape(Nl, H * d))
    return ok_begin, ok_qkv, ok_out


def test_sequence_parallel_exchange_round_trip():
    ret = _run(_seq_job)
    for r in (0, 1):
        assert ret[r] == (True, True, True), (r, ret[r])
