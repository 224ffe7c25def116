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


def test_head_shard_rejects_uneven_split():
    import pytest
    with pytest.raises(ValueError):
        HeadShard(0, 5, 24)
