"""Multi-GPU plumbing: one process per GPU over torch.distributed (NCCL on the GPU
box, gloo in the CPU tests).

Two modes:

* data parallel by image / prompt — the default.  Images are independent (the
  reference is batch-1, single GPU: concept_attention/image_generator.py:135-142), so
  a work list is split round-robin over ranks, every rank runs its own pipeline with
  replicated weights, and the ONLY communication is the final gather of the small
  per-image results.  No collective on the data path.
* head sharding — for very large token counts.  Attention is independent per head
  and the heatmap logits are a plain sum over H*d (concept_attention_pipeline.py:57-61),
  hence over heads; softmax over concepts is not linear so it comes after the sum.
  `HeadShard` describes a rank's head slice, shards the linears tensor-parallel around it
  (column-parallel qkv / mlp.0, row-parallel proj / mlp.2 / linear2) and carries the
  collective the path needs: all-reduce(SUM) of partial heatmap logits and of the
  row-parallel partial outputs (all-gather only when the caller wants the raw vectors).
  `HeadShard(sequence_parallel=True)` is the second form of this mode: weights replicated,
  every token-wise op on 1/world of the rows, heads split only inside the attention through two
  all-to-alls per block (4x fewer NVLink bytes than the tensor-parallel all-reduces and no
  replicated elementwise work).
"""
from __future__ import annotations

import os
from dataclasses import dataclass

import torch
import torch.distributed as dist


def init_distributed(backend: str | None = None):
    """Initialise the default process group from the torchrun environment.
    Returns (rank, world_size, local_rank).  Single process: no group is created."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        kw = {}
        if backend == "nccl":
            torch.cuda.set_device(local_rank)
            kw["device_id"] = torch.device("cuda", local_rank)
        dist.init_process_group(backend, rank=rank, world_size=world, **kw)
    return rank, world, local_rank


def shard_indices(n_items: int, rank: int, world: int) -> list[int]:
    """Round-robin partition of range(n_items): item i goes to rank i % world."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    return list(range(rank, n_items, world))


def gather_results(local: torch.Tensor, n_items: int, rank: int, world: int, dst: int = 0,
                   group=None):
    """Collect per-item results of a round-robin sharded run on `dst`.

    local: [len(shard_indices(n_items, rank, world)), ...] on any device.  Returns
    [n_items, ...] in the ORIGINAL item order on rank `dst` (None elsewhere).  Shards
    are padded to ceil(n_items / world) so one fixed-size gather suffices."""
    if world == 1:
        return local
    per = (n_items + world - 1) // world
    mine = len(shard_indices(n_items, rank, world))
    if local.shape[0] != mine:
        raise ValueError(f"rank {rank}: expected {mine} local results, got {local.shape[0]}")
    pad = torch.zeros((per,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[:mine] = local
    bufs = [torch.empty_like(pad) for _ in range(world)] if rank == dst else None
    dist.gather(pad, bufs, dst=dst, group=group)
    if rank != dst:
        return None
    out = torch.empty((n_items,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    for r in range(world):
        idx = shard_indices(n_items, r, world)
        if idx:
            out[idx] = bufs[r][:len(idx)]
    return out


@dataclass
class HeadShard:
    """Head slice [h0, h1) of rank `rank` among `world` ranks plus the collectives of
    the head-sharded path.  `group=None` uses the default process group."""
    rank: int
    world: int
    num_heads: int
    group: object = None
    reduce_fp32: bool = False     # row-parallel partial sums all-reduced in fp32 (tight parity mode)
    # sequence-parallel form of the head-sharded mode: token-wise ops on 1/world of the rows with
    # REPLICATED weights, heads split only inside the attention (two all-to-alls per block instead
    # of an all-reduce after every row-parallel linear).  B = 1, rows divisible by world.
    sequence_parallel: bool = False
    # sequence-parallel exchanges through peer memory instead of NCCL all-to-all: None = use it
    # when torch's symmetric memory can be set up on this group (NCCL, CUDA), False = never
    peer_exchange: object = None

    def __post_init__(self):
        self._symm = None
        if self.num_heads % self.world:
            raise ValueError(f"{self.num_heads} heads do not split over {self.world} ranks")
        self.h_local = self.num_heads // self.world
        self.h0 = self.rank * self.h_local
        self.h1 = self.h0 + self.h_local

    # -- weight slicing -------------------------------------------------------------
    def qkv_rows(self, head_dim: int = 128) -> torch.Tensor:
        """Row indices of a fused qkv weight [(K H D), in] that produce this rank's heads,
        ordered (K, H_local, D) so the sliced GEMM output is again '(K H D)'."""
        D = self.num_heads * head_dim
        cols = torch.arange(self.h0 * head_dim, self.h1 * head_dim)
        return torch.cat([cols + k * D for k in range(3)])

    def shard_state_dict(self, sd: dict, params) -> dict:
        """Tensor-parallel layout of the head-sharded mode (SURVEY.md 8e "shard linears
        TP-style"), applied IN PLACE to an engine state dict (single-block linear1/linear2
        already split into linear1q|linear1m and linear2a|linear2b):
        qkv / linear1q keep the rows of this rank's heads and mlp.0 / linear1m a 1/G row slice
        (column-parallel, no communication); proj / mlp.2 / linear2 keep the matching input
        columns (row-parallel: partial [B,N,D] sums, one all-reduce each).  The bias of a
        row-parallel linear is added on rank 0 only (zeroed elsewhere)."""
        D = params.hidden_size
        M = int(D * params.mlp_ratio)
        if M % self.world:
            raise ValueError(f"mlp width {M} does not split over {self.world} ranks")
        d = D // self.num_heads
        hcols = slice(self.h0 * d, self.h1 * d)
        mcols = slice(self.rank * (M // self.world), (self.rank + 1) * (M // self.world))

        def col_parallel(prefix, sel):
            for kind in ("weight", "bias"):
                key = f"{prefix}.{kind}"
                if key in sd:
                    idx = sel.to(sd[key].device) if isinstance(sel, torch.Tensor) else sel
                    sd[key] = sd[key][idx].contiguous()

        def row_parallel(prefix, sel, bias_key=None):
            sd[prefix + ".weight"] = sd[prefix + ".weight"][:, sel].contiguous()
            bias_key = bias_key or prefix + ".bias"
            if self.rank > 0 and bias_key in sd:
                sd[bias_key] = torch.zeros_like(sd[bias_key])

        rows = self.qkv_rows(d)
        for i in range(params.depth):
            for st in ("img", "txt"):
                pre = f"double_blocks.{i}.{st}"
                col_parallel(pre + "_attn.qkv", rows)
                row_parallel(pre + "_attn.proj", hcols)
                col_parallel(pre + "_mlp.0", mcols)
                row_parallel(pre + "_mlp.2", mcols)
        for i in range(params.depth_single_blocks):
            pre = f"single_blocks.{i}"
            col_parallel(pre + ".linear1q", rows)
            col_parallel(pre + ".linear1m", mcols)
            row_parallel(pre + ".linear2a", hcols, bias_key=pre + ".linear2.bias")
            sd[pre + ".linear2b.weight"] = sd[pre + ".linear2b.weight"][:, mcols].contiguous()
        return sd

    # -- peer memory ------------------------------------------------------------------
    def symmetric_workspace(self, nbytes: int, device):
        """One symmetric allocation of >= nbytes on every rank of the group, mapped into every
        peer (torch.distributed._symmetric_memory: CUDA VMM handles exchanged at rendezvous).
        Returns (uint8 tensor, handle) — handle.buffer_ptrs_dev is the device array of the W peer
        base pointers, handle.barrier() a cross-GPU barrier kernel on the current stream — or
        None when peer exchange is off / unavailable (gloo, one rank).  COLLECTIVE: every rank
        must call it with the same size at the same point."""
        if self.peer_exchange is False or self.world == 1 or self._gloo():
            return None
        if self._symm is not None and self._symm[0].numel() >= nbytes:
            return self._symm
        try:
            import torch.distributed._symmetric_memory as symm
            grp = self.group if self.group is not None else dist.group.WORLD
            buf = symm.empty(int(nbytes), dtype=torch.uint8, device=device)
            hdl = symm.rendezvous(buf, grp)
            ok = torch.ones(1, device=device)
        except Exception as e:                                  # noqa: BLE001
            if self.peer_exchange is True:
                raise
            import warnings
            warnings.warn(f"peer exchange unavailable ({type(e).__name__}: {e}); using NCCL all-to-all")
            buf = hdl = None
            ok = torch.zeros(1, device=device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.group)   # all ranks or none
        if ok.item() < 1:
            self.peer_exchange = False
            return None
        self._symm = (buf, hdl)
        return self._symm

    # -- collectives ------------------------------------------------------------------
    def _gloo(self) -> bool:
        """gloo (CPU tests, and the 2-process single-GPU test) has no CUDA / flat all_gather:
        use the list form and stage CUDA tensors through the host.  NCCL runs on device."""
        return dist.get_backend(self.group) == "gloo"

    def graph_safe(self) -> bool:
        """True when the collectives can be captured into a CUDA graph (NCCL on device; the gloo
        path stages through the host)."""
        return self.world == 1 or not self._gloo()

    def all_gather(self, x: torch.Tensor, dim: int) -> torch.Tensor:
        """Concatenate the ranks' `x` along `dim` in rank order.  Bit-exact: pure data movement."""
        if self.world == 1:
            return x
        x = x.contiguous()
        if self._gloo():
            parts = [torch.empty(x.shape, dtype=x.dtype) for _ in range(self.world)]
            dist.all_gather(parts, x.cpu(), group=self.group)
            return torch.cat(parts, dim=dim).to(x.device)
        out = torch.empty((self.world,) + tuple(x.shape), dtype=x.dtype, device=x.device)
        dist.all_gather_into_tensor(out, x, group=self.group)
        nd = x.dim()
        dim = dim % nd
        # [G, d0.., dim, ..] -> [d0.., G, dim, ..] -> merge (G, dim)
        perm = list(range(1, dim + 1)) + [0] + list(range(dim + 1, nd + 1))
        shape = list(x.shape)
        shape[dim] *= self.world
        return out.permute(*perm).reshape(shape)

    def all_to_all(self, x: torch.Tensor) -> torch.Tensor:
        """x [world, ...] contiguous: chunk w goes to rank w; returns [world, ...] with chunk w
        received from rank w.  Bit-exact: pure data movement."""
        if self.world == 1:
            return x
        assert x.is_contiguous() and x.shape[0] == self.world
        if self._gloo():
            h = x.cpu().view(torch.uint8)                  # bytes: gloo has no bf16 all-to-all
            out = torch.empty_like(h)
            dist.all_to_all_single(out, h, group=self.group)
            return out.view(x.dtype).view(x.shape).to(x.device)
        out = torch.empty_like(x)
        dist.all_to_all_single(out, x, group=self.group)
        return out

    def all_to_all_begin(self, x: torch.Tensor):
        """Start `all_to_all(x)` and return (out, wait): on NCCL the exchange runs on the process
        group's own stream from here on, so whatever the caller launches before calling `wait()`
        overlaps it; `wait()` makes the current stream wait for the data.  gloo / one rank: done
        synchronously."""
        if self.world == 1 or self._gloo():
            return self.all_to_all(x), (lambda: None)
        assert x.is_contiguous() and x.shape[0] == self.world
        out = torch.empty_like(x)
        work = dist.all_to_all_single(out, x, group=self.group, async_op=True)
        return out, work.wait

    def all_gather_heads(self, x: torch.Tensor) -> torch.Tensor:
        """x [..., H_local*d] (this rank's heads) -> [..., H*d] with heads in rank order."""
        return self.all_gather(x, -1)

    def all_reduce_sum(self, x: torch.Tensor) -> torch.Tensor:
        """In-place SUM over ranks (partial heatmap logits, fp32)."""
        if self.world > 1:
            if self._gloo() and (x.is_cuda or x.dtype == torch.bfloat16):
                h = x.float().cpu()               # gloo: host staging, fp32 arithmetic
                dist.all_reduce(h, op=dist.ReduceOp.SUM, group=self.group)
                x.copy_(h)
            else:
                dist.all_reduce(x, op=dist.ReduceOp.SUM, group=self.group)
        return x
