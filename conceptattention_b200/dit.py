"""B200 engine for the concept-aware Flux DiT.

Host-side mirror of `ModifiedFluxDiT` / `ModifiedDoubleStreamBlock`
(concept_attention/modified_flux_dit.py:29-163,
concept_attention/modified_double_stream_block.py:43-204) with the same forward
contract and state-dict keys, re-planned for one B200:

* tokens of one image live in two buffers, `img [B,P,D]` and `txtc [B,T+C,D]`
  (text rows then concept rows).  Concepts share every txt-stream weight, so
  each txt GEMM also carries the C concept rows and the weight is read once;
  only the modulation vector differs per row segment (K1 kernels select it).
* the big linears are cuBLAS calls through torch (library GEMMs, not claimed);
  everything between them is one C-ABI kernel: LN+modulate (K1a), split +
  QK-RMSNorm + RoPE + layout (K1b), one-way joint attention (K2, tcgen05),
  heatmap logits+softmax+accumulate (K3/K4), gated residual + LN (K1c).
* the reference's four [Nt,L,B,...] vector stacks are never materialised: a
  `HeatmapAccumulator` consumes each selected layer's vectors in place.
"""
from __future__ import annotations

import hashlib
import math

import torch
import torch.nn.functional as F

from . import _cabi, ops
from .weights import FluxParams

BF16 = torch.bfloat16
VECTOR_KEYS = ("output_space_concept_vectors", "output_space_image_vectors",
               "cross_attention_concept_vectors", "cross_attention_image_vectors")


def attention_mode(joint_attention_kwargs) -> int:
    """joint_attention_kwargs -> CA_ATTN_* (modified_double_stream_block.py:126-159)."""
    if joint_attention_kwargs is None:
        return _cabi.ATTN_BOTH
    cross = joint_attention_kwargs.get("concept_cross_attention", True)
    self_ = joint_attention_kwargs.get("concept_self_attention", True)
    if cross and self_:
        return _cabi.ATTN_BOTH
    if cross:
        return _cabi.ATTN_CROSS_ONLY
    if self_:
        return _cabi.ATTN_SELF_ONLY
    return _cabi.ATTN_NEITHER


class HeatmapAccumulator:
    """Running form of compute_heatmaps_from_vectors
    (concept_attention/concept_attention_pipeline.py:29-91).

    softmax (or sparsemax / entmax15) is applied per (timestep, layer) and the
    results are averaged over the selected pairs, exactly as :64-82, but pair by
    pair as the blocks run.  State: fp32 [2, B, C, P] (output space, cross space).
    """

    def __init__(self, B, C, P, device, layer_indices, timesteps=None, softmax=True,
                 attention_norm="sparsemax", round_bf16=False, per_pair=False, buffer=None,
                 pair_buffer=None, pair_slot=None):
        """`per_pair=True` (BASELINE config 4) also keeps every selected pair's normalised map —
        the reference's `heatmaps[timesteps][:, layer_indices]` before the mean
        (concept_attention_pipeline.py:76-77) — in ONE fp32 buffer
        [Nt_sel, L_sel, 2, B, C, P] (unique sorted timesteps x layers; spaces output, cross),
        written by the same kernel launch that updates the running sum.  The slot of the current
        timestep lives in a device int, so a captured forward serves every timestep."""
        if not (softmax or attention_norm in ("softmax", "entmax15", "sparsemax")):
            raise ValueError(f"Unknown attention_norm={attention_norm}")
        self.norm = "softmax" if (softmax or attention_norm == "softmax") else attention_norm
        self.layers = [int(i) for i in layer_indices]
        self.timesteps = None if timesteps is None else [int(t) for t in timesteps]
        self.round_bf16 = round_bf16
        self.B, self.C, self.P = B, C, P
        if buffer is None:
            self.acc = torch.zeros(2, B, C, P, device=device, dtype=torch.float32)
        else:          # caller-owned state (stable address: lets a captured step be replayed)
            assert buffer.shape == (2, B, C, P) and buffer.dtype == torch.float32
            self.acc = buffer.zero_()
        self.count = 0
        self.step = 0
        self.pairs = None
        if per_pair:
            self.pair_steps = sorted(set(self.timesteps)) if self.timesteps is not None else [0]
            self.pair_layers = sorted(set(self.layers))
            shape = (len(self.pair_steps), len(self.pair_layers), 2, B, C, P)
            if pair_buffer is None:
                self.pairs = torch.zeros(shape, device=device, dtype=torch.float32)
            else:
                assert tuple(pair_buffer.shape) == shape and pair_buffer.dtype == torch.float32
                self.pairs = pair_buffer.zero_()
            # caller-owned like the buffers: a captured forward bakes this address in
            self.pair_slot = (torch.zeros(1, device=device, dtype=torch.int32) if pair_slot is None
                              else pair_slot.zero_())
            assert self.pair_slot.dtype == torch.int32 and self.pair_slot.numel() == 1

    @property
    def per_pair(self):
        return self.pairs

    def begin_step(self, step_index: int):
        self.step = step_index
        if self.pairs is not None and step_index in self.pair_steps:
            # first slot of this timestep; read by the kernels at run time (graph-replay safe)
            self.pair_slot.fill_(self.pair_steps.index(step_index) * len(self.pair_layers) * 2)

    def multiplicity(self, layer: int) -> int:
        mt = 1 if self.timesteps is None else self.timesteps.count(self.step)
        return mt * self.layers.count(layer)

    def _pair_args(self, layer, space):
        if self.pairs is None:
            return {}
        return dict(pair=self.pairs.view(-1, self.B, self.C, self.P), pair_slot=self.pair_slot,
                    pair_slot_off=self.pair_layers.index(layer) * 2 + space)

    def add(self, layer, attn_img, attn_con, q_pre, P, T):
        """attn_img [B,P,D]; attn_con [B,C,D] view; q_pre [B,H,P+T+C,128]."""
        m = self.multiplicity(layer)
        for rep in range(m):
            # output space and cross-attention space of this layer in ONE launch
            kw = {}
            if rep == 0 and self.pairs is not None:
                a0, a1 = self._pair_args(layer, 0), self._pair_args(layer, 1)
                kw = dict(pair=a0["pair"], pair_slot=a0["pair_slot"],
                          pair_slot_off=(a0["pair_slot_off"], a1["pair_slot_off"]))
            ops.heatmap_accumulate_both(attn_img, attn_con, self.acc[0], q_pre[:, :, :P], q_pre[:, :, P + T:],
                                        self.acc[1], self.norm, self.round_bf16, **kw)
        self.count += m

    def signature(self, depth):
        """What a captured forward bakes in for this accumulator at the current step."""
        return (self.acc.data_ptr(), self.norm, self.round_bf16,
                tuple(self.multiplicity(i) for i in range(depth)),
                None if self.pairs is None else (self.pairs.data_ptr(), self.pair_slot.data_ptr(),
                                                 tuple(self.pair_layers)))

    def note_replay(self, depth):
        """Host bookkeeping of one replayed forward (the kernels ran inside the graph)."""
        self.count += sum(self.multiplicity(i) for i in range(depth))

    def add_logits(self, layer, logits):
        """Head-sharded form: `logits` fp32 [2,B,C,P] (output space, cross space) are the
        all-reduced sums of the ranks' partial logits; normalise + accumulate (K4)."""
        m = self.multiplicity(layer)
        for rep in range(m):
            for s in range(2):
                pair = None
                if rep == 0 and self.pairs is not None:       # slot chosen on the host here
                    pair = self.pairs[self.pair_steps.index(self.step), self.pair_layers.index(layer), s]
                ops.norm_accumulate(logits[s], self.acc[s], self.norm, self.round_bf16, pair=pair)
        self.count += m

    def per_pair_heatmaps(self):
        """-> (concept [Nt_sel,L_sel,B,C,h,w], cross [...]) fp32 views of the per-pair buffer,
        axes ordered as `pair_steps` x `pair_layers`."""
        if self.pairs is None:
            raise ValueError("accumulator was built without per_pair=True")
        side = math.isqrt(self.P)
        shp = self.pairs.shape[:2] + ((self.B, self.C, side, side) if side * side == self.P
                                      else (self.B, self.C, self.P))
        return self.pairs[:, :, 0].reshape(shp), self.pairs[:, :, 1].reshape(shp)

    def finalize(self):
        """-> (concept_heatmaps, cross_attention_maps), each [B,C,h,w] fp32."""
        if self.count == 0:
            raise ValueError("no (timestep, layer) pair was selected")
        out = ops.finalize_heatmaps(self.acc.view(2 * self.B, self.C, self.P), self.count,
                                    self.round_bf16)
        out = out.view(2, self.B, self.C, *out.shape[2:])
        return out[0], out[1]


def timestep_embedding(t, dim, max_period=10000, time_factor=1000.0):
    """flux/src/flux/modules/layers.py:28-49 (cos | sin halves, fp32 then cast)."""
    t = time_factor * t
    half = dim // 2
    freqs = torch.exp(-math.log(max_period) *
                      torch.arange(half, dtype=torch.float32, device=t.device) / half)
    args = t[:, None].float() * freqs[None]
    emb = torch.cat([torch.cos(args), torch.sin(args)], dim=-1)
    return emb.to(t) if torch.is_floating_point(t) else emb


def rope_cos_sin(ids, axes_dim, theta):
    """ids [N,A] -> fp32 [N, sum(axes)/2, 2] (cos, sin); angles in fp64 like
    flux/src/flux/math.py:15-22."""
    parts = []
    for a, dim in enumerate(axes_dim):
        expo = torch.arange(0, dim, 2, dtype=torch.float64, device=ids.device) / dim
        omega = 1.0 / (float(theta) ** expo)
        parts.append(ids[:, a].to(torch.float64)[:, None] * omega[None, :])
    ang = torch.cat(parts, dim=1)
    return torch.stack([torch.cos(ang), torch.sin(ang)], dim=-1).float().contiguous()


class FluxDiTEngine:
    """Concept-aware Flux DiT on one B200.  `state_dict` uses the reference keys
    (see weights.linear_shapes); tensors must be bf16 on `device`."""

    def __init__(self, params: FluxParams, state_dict: dict, device="cuda:0", head_shard=None):
        """`head_shard` (parallel.HeadShard, one image over several GPUs):
        tensor-parallel form - this rank's head slice of every attention (qkv rows, K1b, K2, K3
        partial logits), linears sharded around it, partial sums and partial heatmap logits
        all-reduced; `sequence_parallel=True` - whole weights, this rank's 1/world of the rows
        for every token-wise op, heads split only inside the attention (two all-to-alls)."""
        self.p = params
        self.device = torch.device(device)
        self.hs = head_shard
        lib = _cabi.load()
        idx = self.device.index if self.device.index is not None else torch.cuda.current_device()
        _cabi.check(lib.ca_device_check(idx), "ca_device_check")
        D = params.hidden_size
        if D // params.num_heads != 128 or sum(params.axes_dim) != 128:
            raise ValueError("the sm_100a attention kernel is built for head_dim 128")
        self.sd = {}
        for k, v in state_dict.items():
            v = v.to(device=self.device, dtype=BF16).contiguous()
            self.sd[k] = v
        # linear2 of the single blocks consumes cat(attn, gelu(mlp)); keep its two
        # column blocks as separate contiguous matrices so no cat is ever built.
        for i in range(params.depth_single_blocks):
            w = self.sd.pop(f"single_blocks.{i}.linear2.weight")
            self.sd[f"single_blocks.{i}.linear2a.weight"] = w[:, :D].contiguous()
            self.sd[f"single_blocks.{i}.linear2b.weight"] = w[:, D:].contiguous()
            # linear1 = [qkv | mlp] rows: two GEMMs so the mlp half can carry the GELU epilogue
            w1, b1 = self.sd.pop(f"single_blocks.{i}.linear1.weight"), self.sd.pop(f"single_blocks.{i}.linear1.bias")
            self.sd[f"single_blocks.{i}.linear1q.weight"] = w1[:3 * D].contiguous()
            self.sd[f"single_blocks.{i}.linear1q.bias"] = b1[:3 * D].contiguous()
            self.sd[f"single_blocks.{i}.linear1m.weight"] = w1[3 * D:].contiguous()
            self.sd[f"single_blocks.{i}.linear1m.bias"] = b1[3 * D:].contiguous()
        # every modulation vector of a forward depends only on (vec, concept_vec): ONE
        # [2B,D]x[D,Mtot] GEMM per forward instead of 3 per double + 1 per single block
        # (layers.py:113-126 applied 96 times).  The originals become views of the stack.
        names = [f"double_blocks.{i}.{s}_mod.lin" for i in range(params.depth) for s in ("img", "txt")]
        names += [f"single_blocks.{i}.modulation.lin" for i in range(params.depth_single_blocks)]
        names += ["final_layer.adaLN_modulation.1"]
        self.mod_off, off = {}, 0
        for n in names:
            self.mod_off[n] = (off, self.sd[n + ".weight"].shape[0])
            off += self.sd[n + ".weight"].shape[0]
        self.mod_w = torch.empty(off, D, device=self.device, dtype=BF16)
        self.mod_b = torch.empty(off, device=self.device, dtype=BF16)
        for n in names:
            o, m = self.mod_off[n]
            self.mod_w[o:o + m].copy_(self.sd.pop(n + ".weight"))
            self.mod_b[o:o + m].copy_(self.sd.pop(n + ".bias"))
        if self.hs is not None and self.hs.world > 1 and not self.hs.sequence_parallel:
            self.hs.shard_state_dict(self.sd, params)       # sequence-parallel: weights stay whole
        self._rope_cache = {}
        # CUDA graphs of whole forwards (one per signature), captured the second time a
        # signature is seen; the first call runs launch by launch and warms cuBLAS.
        self.use_graphs = True
        # the text+concept stream of a double block (five small, latency-bound launches per
        # GEMM pair) runs on a second CUDA stream beside the image stream; captured graphs keep
        # the fork/join as parallel branches.  Off in the tensor-parallel head-sharded mode (one NCCL
        # communicator must see its collectives from one stream).
        self.use_side_stream = (head_shard is None or head_shard.world == 1
                                or head_shard.sequence_parallel)   # its text chain has no collective
        self.skip_unused_tail = True      # see the double-block loop (encode path only)
        # head-sharded mode: two token chunks on two streams so one chunk computes while the other's
        # partial sums are in the all-reduce.  Measured on 8 GPUs (config 5): 315 -> 353 ms per image
        # (twice the launches; the ranks are launch-bound at 2 host cores each) -> off by default.
        self.overlap_allreduce = False
        self.tp_chunk_min_rows = 2048
        self._tp_side = None
        self._side = None
        self.max_graphs = 8
        self._graphs, self._graph_seen, self._graph_pool = {}, {}, None

    def _lin_rp(self, prefix, x, bias_key=None, extra=None):
        """Row-parallel linear: y = x W^T (+ xe We^T) + b summed over the ranks; a plain linear
        on one GPU.  `extra` = (xe, We) is a second accumulating GEMM (linear2 over
        cat(attn, gelu(mlp)) without building the cat).  Sharded: W, We hold this rank's input
        columns and b is zero on ranks > 0.  `hs.reduce_fp32` keeps the partial sums in fp32
        through the all-reduce (one bf16 rounding, as on one GPU) at twice the NVLink bytes."""
        w, b = self.sd[prefix + ".weight"], self.sd.get(bias_key or prefix + ".bias")
        sharded = self.hs is not None and self.hs.world > 1 and not self.hs.sequence_parallel
        x2 = x.reshape(-1, x.shape[-1])
        if sharded and self.hs.reduce_fp32:
            y = torch.mm(x2, w.t(), out_dtype=torch.float32)
            if extra is not None:
                y += torch.mm(extra[0].reshape(-1, extra[0].shape[-1]), extra[1].t(), out_dtype=torch.float32)
            if b is not None and self.hs.rank == 0:
                y += b
            return self.hs.all_reduce_sum(y).to(BF16).view(*x.shape[:-1], w.shape[0])
        y = F.linear(x2, w, b)
        if extra is not None:
            y.addmm_(extra[0].reshape(-1, extra[0].shape[-1]), extra[1].t())
        if sharded:
            self.hs.all_reduce_sum(y)
        return y.view(*x.shape[:-1], w.shape[0])

    # -- helpers ---------------------------------------------------------------------
    def _lin(self, prefix, x):
        return F.linear(x, self.sd[prefix + ".weight"], self.sd.get(prefix + ".bias"))

    def _lin_gelu(self, prefix, x):
        """gelu_tanh(x W^T + b) with the activation in the cuBLASLt epilogue (library GEMM)."""
        w, b = self.sd[prefix + ".weight"], self.sd[prefix + ".bias"]
        y = torch._addmm_activation(b, x.reshape(-1, x.shape[-1]), w.t(), use_gelu=True)
        return y.view(*x.shape[:-1], w.shape[0])

    def _mlp_embed(self, prefix, x):
        return self._lin(prefix + ".out_layer", F.silu(self._lin(prefix + ".in_layer", x)))

    def _rope(self, img_ids, txt_ids, concept_ids):
        """cos/sin table for the packed order [img | txt | concept]; ids are taken
        from batch element 0 (prepare()/embed_concepts() repeat them per batch).
        Cached per id tensors so a denoising loop builds it once."""
        parts = (img_ids[0], txt_ids[0], concept_ids[0])
        def ver(t):                       # inference tensors have no version counter
            try:
                return t._version
            except RuntimeError:
                return -1
        key = tuple((t.data_ptr(), tuple(t.shape), ver(t), str(t.device)) for t in parts)
        if self._rope_cache.get("key") != key:
            # new id tensors: look the table up by CONTENT (one small D2H per new id set) so
            # equal ids always map to the same table address (captured graphs bake it in)
            for full, name in ((img_ids, "img_ids"), (txt_ids, "txt_ids"), (concept_ids, "concept_ids")):
                if full.shape[0] > 1 and not bool((full == full[:1]).all()):
                    raise ValueError(f"{name} differ across the batch: one RoPE table serves the whole batch")
            ids = torch.cat([t.to(self.device, torch.float32) for t in parts], 0)
            digest = hashlib.sha1(ids.cpu().numpy().tobytes()).hexdigest() + str(tuple(ids.shape))
            tables = self._rope_cache.setdefault("tables", {})
            if digest not in tables:
                if len(tables) >= 16:
                    tables.pop(next(iter(tables)))
                tables[digest] = rope_cos_sin(ids, self.p.axes_dim, self.p.theta)
            # keep the id tensors alive so data_ptr cannot be recycled under the key
            self._rope_cache.update(key=key, ids=parts, table=tables[digest])
        return self._rope_cache["table"]

    def _peer_regions(self, Pl, Tl, C, dev):
        """Symmetric-memory regions of the fused sequence-parallel exchange (None: NCCL path):
        send [W,3,Hl,Pl+Tl+C,128] | K2 out_img [W*Pl,Hl*128] | K2 out_txtc [W*Tl+C,Hl*128], bf16,
        sized for the double blocks (single blocks use the same offsets with C = 0)."""
        hs = self.hs
        W, Hl = hs.world, hs.h_local
        al = lambda n: (n + 1023) // 1024 * 1024
        send_b = al(W * 3 * Hl * (Pl + Tl + C) * 256)
        img_b = al(W * Pl * Hl * 256)
        tc_b = al((W * Tl + C) * Hl * 256)
        sym = hs.symmetric_workspace(send_b + img_b + tc_b, dev)
        if sym is None:
            return None
        buf, hdl = sym

        def view(off, shape):
            n = 2
            for d in shape:
                n *= d
            return buf[off:off + n].view(torch.bfloat16).view(shape)
        return {"hdl": hdl, "ptrs": hdl.buffer_ptrs_dev, "view": view,
                "send_off": 0, "img_off": send_b, "tc_off": send_b + img_b}

    def _rope_local(self, rope, P, T):
        """Sequence-parallel mode: this rank's rows [img_l | txt_l | concept] of the full table
        (rows [img P | txt T | concept]).  One table per (full table, P, T), kept beside the full
        table it was cut from, so its address is as stable as the full table's (captured graphs
        bake it in; `_capture` also holds a reference)."""
        W, rk = self.hs.world, self.hs.rank
        key = (rope.data_ptr(), P, T)
        local = self._rope_cache.setdefault("locals", {})
        if key not in local:
            if len(local) >= 16:
                local.pop(next(iter(local)))
            Pl, Tl = P // W, T // W
            local[key] = (rope, torch.cat((rope[rk * Pl:(rk + 1) * Pl], rope[P + rk * Tl:P + (rk + 1) * Tl],
                                           rope[P + T:]), 0).contiguous())
        return local[key][1]

    # -- forward -----------------------------------------------------------------------
    @torch.no_grad()
    def forward(self, img, img_ids, txt, txt_ids, concepts, concept_ids, concept_vec, timesteps, y,
                guidance=None, stop_after_multimodal_attentions=False, edit_metadata=None,
                iteration=None, joint_attention_kwargs=None, accumulator: HeatmapAccumulator = None,
                collect_vectors=True, **kwargs):
        """ModifiedFluxDiT.forward contract (see `_forward`).  The fast path
        (`collect_vectors=False`, one GPU, running accumulator) replays a CUDA graph of the
        whole forward: ~750 launches per timestep become one cudaGraphLaunch."""
        args = dict(img=img, txt=txt, concepts=concepts, concept_vec=concept_vec,
                    timesteps=timesteps, y=y, guidance=guidance)
        rest = dict(img_ids=img_ids, txt_ids=txt_ids, concept_ids=concept_ids,
                    stop_after_multimodal_attentions=stop_after_multimodal_attentions,
                    joint_attention_kwargs=joint_attention_kwargs, accumulator=accumulator,
                    collect_vectors=collect_vectors)
        eligible = (self.use_graphs and not collect_vectors and (self.hs is None or self.hs.graph_safe())
                    and ops.ATTN_EVENTS is None and ops.HBM_EVENTS is None
                    and (accumulator is None or accumulator.per_pair is None or self.hs is None
                         or self.hs.world == 1)
                    and concept_vec is not None and img.ndim == 3 and txt.ndim == 3
                    and not torch.cuda.is_current_stream_capturing())
        if not eligible:
            return self._forward(**args, **rest)
        if self.p.guidance_embed and guidance is None:
            raise ValueError("Didn't get guidance strength for guidance distilled model.")
        if not self.p.guidance_embed:
            args["guidance"] = None
        rope = self._rope(img_ids, txt_ids, concept_ids)
        seqpar = self.hs is not None and self.hs.world > 1 and self.hs.sequence_parallel
        if seqpar and img.shape[1] % self.hs.world == 0 and txt.shape[1] % self.hs.world == 0:
            # cut this rank's rows of the table OUTSIDE any capture (stable address, eager memory)
            self._rope_local(rope, img.shape[1], txt.shape[1])
        key = (tuple((k, tuple(v.shape)) for k, v in args.items() if v is not None),
               attention_mode(joint_attention_kwargs), bool(stop_after_multimodal_attentions),
               rope.data_ptr(), accumulator.signature(self.p.depth) if accumulator is not None else None)
        ent = self._graphs.get(key)
        if ent is None:
            self._graph_seen[key] = self._graph_seen.get(key, 0) + 1
            if self._graph_seen[key] < 2:
                return self._forward(**args, **rest)
            if len(self._graphs) >= self.max_graphs:          # drop the oldest capture
                self._graph_seen.pop(next(iter(self._graphs)), None)
                self._graphs.pop(next(iter(self._graphs)))
            if len(self._graph_seen) > 4 * self.max_graphs:   # signatures seen once and never again
                for k in [k for k in self._graph_seen if k not in self._graphs][:2 * self.max_graphs]:
                    del self._graph_seen[k]
            ent = self._capture(args, rest, rope)
            self._graphs[key] = ent
        elif accumulator is not None:
            accumulator.note_replay(self.p.depth)
        for k, v in args.items():
            if v is not None:
                ent["static"][k].copy_(v, non_blocking=True)
        ent["graph"].replay()
        ops.LAUNCHES += ent["n_launch"]
        # a fresh tensor per call, like the launch-by-launch path (the graph's own output buffer is
        # overwritten by the next replay of this signature)
        return (ent["pred"].clone() if ent["pred"] is not None else None), {}

    def _capture(self, args, rest, rope):
        """Capture one forward on static copies of the per-call tensors.  The accumulator's
        host bookkeeping (count) happens once here, so the caller replays right after."""
        dev = self.device
        static = {k: torch.empty(v.shape, device=dev, dtype=BF16) for k, v in args.items() if v is not None}
        torch.cuda.synchronize(dev)
        graph = torch.cuda.CUDAGraph()
        n0 = ops.LAUNCHES
        if self._graph_pool is None:
            self._graph_pool = torch.cuda.graph_pool_handle()
        with torch.cuda.graph(graph, pool=self._graph_pool):
            pred, _ = self._forward(**{k: static.get(k) for k in args}, **rest)
        acc = rest["accumulator"]
        local = self._rope_cache.get("locals", {}).get((rope.data_ptr(), args["img"].shape[1], args["txt"].shape[1]))
        return {"graph": graph, "static": static, "pred": pred, "n_launch": ops.LAUNCHES - n0,
                "keep": (rope, local, None if acc is None else (acc.acc, acc.pairs, getattr(acc, "pair_slot", None)))}

    @torch.no_grad()
    def _forward(self, img, img_ids, txt, txt_ids, concepts, concept_ids, concept_vec, timesteps, y,
                guidance=None, stop_after_multimodal_attentions=False, edit_metadata=None,
                iteration=None, joint_attention_kwargs=None, accumulator: HeatmapAccumulator = None,
                collect_vectors=True, **kwargs):
        """Same contract as ModifiedFluxDiT.forward (modified_flux_dit.py:75-163):
        returns (pred | None, {key: [L,B,...]}).  With `collect_vectors=False` the
        dict is empty and heatmaps only flow into `accumulator` (the fast path)."""
        assert concept_vec is not None, "Concept vectors must be provided for this implementation."
        if img.ndim != 3 or txt.ndim != 3:
            raise ValueError("Input img and txt tensors must have 3 dimensions.")
        p, dev = self.p, self.device
        D, H = p.hidden_size, p.num_heads
        to = lambda t: t.to(device=dev, dtype=BF16)
        img, txt, concepts, y, concept_vec = map(to, (img, txt, concepts, y, concept_vec))
        timesteps = to(timesteps)
        B, P, T, C = img.shape[0], img.shape[1], txt.shape[1], concepts.shape[1]
        mode = attention_mode(joint_attention_kwargs)

        def cond(pooled):
            v = self._mlp_embed("time_in", timestep_embedding(timesteps, 256))
            if p.guidance_embed:
                if guidance is None:
                    raise ValueError("Didn't get guidance strength for guidance distilled model.")
                v = v + self._mlp_embed("guidance_in", timestep_embedding(to(guidance), 256))
            return v + self._mlp_embed("vector_in", pooled)

        rope = self._rope(img_ids, txt_ids, concept_ids)
        sharded = self.hs is not None and self.hs.world > 1
        seqpar = sharded and self.hs.sequence_parallel
        if seqpar:
            # sequence-parallel: from here on P and T are THIS RANK's row counts; only the
            # attention (see `attend`) sees all rows, for this rank's heads
            W, rk = self.hs.world, self.hs.rank
            if B != 1 or P % W or T % W:
                raise ValueError(f"sequence-parallel mode needs B = 1 and P, T divisible by {W} (got {B}, {P}, {T})")
            Pg, Tg, P, T = P, T, P // W, T // W
            img, txt = img[:, rk * P:(rk + 1) * P], txt[:, rk * T:(rk + 1) * T]
            rope = self._rope_local(rope, Pg, Tg)            # rows [img_l | txt_l | concept]
        peer = self._peer_regions(P, T, C, dev) if seqpar else None
        x_img = self._lin("img_in", img).contiguous()
        vec, cvec = cond(y), cond(concept_vec)
        txtc = self._lin("txt_in", torch.cat((txt, concepts), 1)).contiguous()   # [B,T+C,D]
        svec = F.silu(torch.cat((vec, cvec), 0))                                  # [2B,D]
        mods = F.linear(svec, self.mod_w, self.mod_b)                             # [2B,Mtot]

        # every kernel addresses its modulation chunks by ABSOLUTE column offset into `mods`:
        # rows [0,B) come from vec (img / txt / single streams), rows [B,2B) from concept_vec
        m_vec, m_con = mods[:B], mods[B:]
        off = lambda name: self.mod_off[name][0]

        collected = {k: [] for k in VECTOR_KEYS} if collect_vectors else None
        tp = sharded and not seqpar                              # tensor-parallel linears
        Hl = self.hs.h_local if tp else H                        # heads of the qkv GEMMs on this rank
        # LN + modulate of block 0; later blocks get theirs fused into the previous block's last
        # gated-residual kernel (K1c with the NEXT block's shift/scale): one pass over x less per
        # block boundary
        o_i, o_t = off("double_blocks.0.img_mod.lin"), off("double_blocks.0.txt_mod.lin")
        ln = {"img": ops.ln_modulate(x_img, m_vec, None, None, o_i, o_i + D),
              "txt": ops.ln_modulate(txtc, m_vec, m_con, T, o_t, o_t + D)}
        side = None
        if self.use_side_stream and not tp:
            if self._side is None:
                self._side = torch.cuda.Stream(device=dev)
            side = self._side
        main = torch.cuda.current_stream(dev)

        # Head-sharded mode: every row-parallel GEMM is followed by an all-reduce that is several
        # times longer than the GEMM itself.  The post-attention chains are row-wise, so they
        # run as TWO token chunks on two CUDA streams: while one chunk's partial sums are in
        # the NCCL all-reduce, the other chunk computes (GEMMs, K1c).  One GPU: one chunk.
        def chunked(n_rows, fn):
            """fn(r0, r1) over [0, n_rows); two chunks on (main, tp stream) when sharded."""
            if not tp or B != 1 or n_rows < self.tp_chunk_min_rows or not self.overlap_allreduce:
                fn(0, n_rows)
                return
            if self._tp_side is None:
                self._tp_side = torch.cuda.Stream(device=dev)
            mid = (n_rows // 2 + 127) // 128 * 128
            self._tp_side.wait_stream(main)
            with torch.cuda.stream(self._tp_side):
                fn(mid, n_rows)
            fn(0, mid)
            main.wait_stream(self._tp_side)

        def attend(qkv_a, qkv_b, scales_a, scales_b, rope_x, n_con, amode, want_q_pre=False, out=None,
                   meanwhile=None):
            """K1b + K2 -> (attn_img [B,P,*], attn_tc [B,T+n_con,*], q_pre | None) for this rank's
            rows.  Sequence-parallel: K1b writes the all-to-all send buffer, K2 runs this rank's
            heads over ALL rows, a second all-to-all brings back every head of the local rows.
            `meanwhile()` is work that does not depend on the attention (the MLP-in GEMM of a single
            block): it is launched while the q/k/v exchange is on NVLink."""
            if not seqpar:
                if meanwhile is not None:
                    meanwhile()
                res = ops.qkv_prep(qkv_a, qkv_b, scales_a, scales_b, rope_x, Hl, want_q_pre=want_q_pre)
                o = (None, None) if out is None else out
                a_img, a_tc = ops.joint_attention(res[0], res[1], res[2], P, T, n_con, amode, o[0], o[1])
                return a_img, a_tc, (res[3] if want_q_pre else None)
            if peer is not None:
                # fused exchange over peer memory: K1b writes this rank's send region, ONE cross-GPU
                # barrier, then the gather kernel pulls the q/k/v of this rank's heads out of every
                # peer; K2 writes its outputs into the symmetric region, one barrier, and the
                # second gather pulls this rank's rows of every peer's heads.  A region is written
                # again only after the NEXT barrier, which a peer reaches after its reads.
                hloc, nloc = self.hs.h_local, P + T + n_con
                send = peer["view"](peer["send_off"], (W, 3, hloc, nloc, 128))
                o_img = peer["view"](peer["img_off"], (1, W * P, hloc * 128))
                o_tc = peer["view"](peer["tc_off"], (1, W * T + n_con, hloc * 128))
                _, q_pre = ops.qkv_prep_grouped(qkv_a, qkv_b, scales_a, scales_b, rope_x, H, hloc,
                                                want_q_pre=want_q_pre, send=send)
                if meanwhile is not None:
                    meanwhile()
                peer["hdl"].barrier(channel=0)
                q, k, v = ops.seq_gather_qkv(peer["ptrs"], peer["send_off"], W, rk, hloc, P, T, n_con, dev)
                ops.joint_attention(q, k, v, Pg, Tg, n_con, amode, o_img, o_tc)
                peer["hdl"].barrier(channel=0)
                a_img, a_tc = ops.seq_gather_out(peer["ptrs"], peer["img_off"], peer["tc_off"], W, rk, hloc,
                                                 P, T, n_con, dev, out)
                return a_img, a_tc, q_pre
            send, q_pre = ops.qkv_prep_grouped(qkv_a, qkv_b, scales_a, scales_b, rope_x, H,
                                               self.hs.h_local, want_q_pre=want_q_pre)
            recv, wait = self.hs.all_to_all_begin(send)
            if meanwhile is not None:
                meanwhile()
            wait()
            q, k, v = ops.seq_unpack_qkv(recv, rk, P, T, n_con)
            o_img, o_tc = ops.joint_attention(q, k, v, Pg, Tg, n_con, amode)
            back = self.hs.all_to_all(ops.seq_pack_out(o_img, o_tc, W, P, T, n_con))
            a_img, a_tc = ops.seq_unpack_out(back, P, T, n_con, out)
            return a_img, a_tc, q_pre

        def txt_chain(i, attn_tc):
            """proj -> gated residual + LN -> MLP -> gated residual (+ next block's LN) -> next
            block's qkv GEMM, all on the text+concept rows; returns the next tc_qkv (or None)."""
            pre = f"double_blocks.{i}"
            o = off(pre + ".txt_mod.lin")
            proj = self._lin_rp(pre + ".txt_attn.proj", attn_tc)
            h = ops.gate_residual_ln_modulate(txtc, proj, m_vec, m_con, T, o + 2 * D, o + 3 * D, o + 4 * D)
            h = self._lin_rp(pre + ".txt_mlp.2", self._lin_gelu(pre + ".txt_mlp.0", h))
            if i + 1 < p.depth:
                on = off(f"double_blocks.{i + 1}.txt_mod.lin")
                ln_t = ops.gate_residual_ln_modulate(txtc, h, m_vec, m_con, T, o + 5 * D, on, on + D)
                return self._lin(f"double_blocks.{i + 1}.txt_attn.qkv", ln_t)
            ops.gate_residual_ln_modulate(txtc, h, m_vec, m_con, T, o + 5 * D, 0, 0, want_out=False)
            return None

        tc_qkv = self._lin("double_blocks.0.txt_attn.qkv", ln.pop("txt")) if p.depth else None
        for i in range(p.depth):
            pre = f"double_blocks.{i}"
            sel = collect_vectors or (accumulator is not None and accumulator.multiplicity(i) > 0)
            img_qkv = self._lin(pre + ".img_attn.qkv", ln["img"])
            if side is not None and i > 0:
                main.wait_stream(side)                   # join: tc_qkv (and txtc) of the side branch
            attn_img, attn_tc, q_pre = attend(
                img_qkv, tc_qkv,
                (self.sd[pre + ".img_attn.norm.query_norm.scale"], self.sd[pre + ".img_attn.norm.key_norm.scale"]),
                (self.sd[pre + ".txt_attn.norm.query_norm.scale"], self.sd[pre + ".txt_attn.norm.key_norm.scale"]),
                rope, C, mode, want_q_pre=sel)
            if sel and accumulator is not None and accumulator.multiplicity(i) > 0:
                if seqpar:
                    # complete logits (all heads) of this rank's pixels, gathered along P (small)
                    part = torch.empty(2, B, C, P, device=dev, dtype=torch.float32)
                    ops.heatmap_logits(attn_img, attn_tc[:, T:], out=part[0])
                    ops.heatmap_logits(q_pre[:, :, :P], q_pre[:, :, P + T:], per_head=True, out=part[1])
                    accumulator.add_logits(i, self.hs.all_gather(part, -1))
                elif tp:
                    # partial logits over this rank's heads, ONE all-reduce for both spaces
                    part = torch.empty(2, B, C, P, device=dev, dtype=torch.float32)
                    ops.heatmap_logits(attn_img, attn_tc[:, T:], out=part[0])
                    ops.heatmap_logits(q_pre[:, :, :P], q_pre[:, :, P + T:], per_head=True, out=part[1])
                    accumulator.add_logits(i, self.hs.all_reduce_sum(part))
                else:
                    accumulator.add(i, attn_img, attn_tc[:, T:], q_pre, P, T)
            if tp and collect_vectors:
                q_pre = self.hs.all_gather(q_pre, 1)
            if sel and collect_vectors and seqpar:         # image rows gathered along the token axis
                collected["output_space_concept_vectors"].append(attn_tc[:, T:].clone())
                collected["output_space_image_vectors"].append(self.hs.all_gather(attn_img, 1))
                collected["cross_attention_concept_vectors"].append(q_pre[:, :, P + T:].clone())
                collected["cross_attention_image_vectors"].append(self.hs.all_gather(q_pre[:, :, :P], 2))
            elif sel and collect_vectors:
                gather = (lambda t: self.hs.all_gather_heads(t)) if tp else (lambda t: t)
                collected["output_space_concept_vectors"].append(gather(attn_tc[:, T:].clone()))
                collected["output_space_image_vectors"].append(gather(attn_img))
                collected["cross_attention_concept_vectors"].append(q_pre[:, :, P + T:].clone())
                collected["cross_attention_image_vectors"].append(q_pre[:, :, :P].clone())
            # encode_image path (stop_after_multimodal_attentions, heatmaps only): nothing after
            # the last selected layer's attention is ever read — the reference still runs those
            # proj/MLP updates (and whole blocks) and discards them
            # (modified_flux_dit.py:147-153); here they are not computed.
            if (self.skip_unused_tail and stop_after_multimodal_attentions and not collect_vectors
                    and accumulator is not None
                    and not any(accumulator.multiplicity(l) > 0 for l in range(i + 1, p.depth))):
                break
            # residual updates.  Text+concept rows (txt weights, per-segment gates) fork onto the
            # side stream right after the attention; the image rows stay on the main stream.  The
            # side branch runs on into the next block's txt qkv GEMM and is joined before K1b.
            if side is not None:
                side.wait_stream(main)
                with torch.cuda.stream(side):
                    tc_qkv = txt_chain(i, attn_tc)
                # (no record_stream needed: every tensor the branch reads outlives the join at the
                # next block's K1b, and everything it allocates is first reused after a later fork)
            else:
                tc_qkv = txt_chain(i, attn_tc)
            o = off(pre + ".img_mod.lin")
            last = i + 1 == p.depth
            on = 0 if last else off(f"double_blocks.{i + 1}.img_mod.lin")
            ln_next = None if last else torch.empty_like(x_img)

            def img_rows(r0, r1, o=o, on=on, last=last, ln_next=ln_next, attn_img=attn_img, pre=pre):
                xr = x_img[:, r0:r1]
                proj = self._lin_rp(pre + ".img_attn.proj", attn_img[:, r0:r1])
                h = ops.gate_residual_ln_modulate(xr, proj, m_vec, None, None, o + 2 * D, o + 3 * D, o + 4 * D)
                h = self._lin_rp(pre + ".img_mlp.2", self._lin_gelu(pre + ".img_mlp.0", h))
                if last:
                    ops.gate_residual_ln_modulate(xr, h, m_vec, None, None, o + 5 * D, 0, 0, want_out=False)
                else:
                    ops.gate_residual_ln_modulate(xr, h, m_vec, None, None, o + 5 * D, on, on + D,
                                                  out=ln_next[:, r0:r1])

            chunked(P, img_rows)
            if not last:
                ln["img"] = ln_next
        if side is not None and p.depth:
            main.wait_stream(side)                       # txtc is final
        if collect_vectors:
            collected = {k: torch.stack(v, 0) for k, v in collected.items()}
        else:
            collected = {}
        if stop_after_multimodal_attentions:
            return None, collected

        # ---- single-stream blocks over [img | txt] (modified_single_stream_block.py:43-56)
        x = torch.cat((x_img, txtc[:, :T]), 1).contiguous()                      # [B,P+T,D]
        N = P + T
        rope_s = rope[:N].contiguous() if C > 0 else rope
        o = off("single_blocks.0.modulation.lin") if p.depth_single_blocks else 0
        xm = ops.ln_modulate(x, m_vec, None, None, o, o + D) if p.depth_single_blocks else None
        for i in range(p.depth_single_blocks):
            pre = f"single_blocks.{i}"
            o = off(pre + ".modulation.lin")
            h = self._lin(pre + ".linear1q", xm)                                  # [B,N,3*Hl*128]
            box = {}

            def mlp_in(box=box, pre=pre, xm=xm):                                  # [B,N,M/G] gelu in epilogue
                box["g"] = self._lin_gelu(pre + ".linear1m", xm)
            scales = (self.sd[pre + ".norm.query_norm.scale"], self.sd[pre + ".norm.key_norm.scale"])
            if B == 1:
                attn = torch.empty(1, N, Hl * 128, device=dev, dtype=BF16)
                attend(h, None, scales, None, rope_s, 0, _cabi.ATTN_BOTH, out=(attn[:, :P], attn[:, P:]),
                       meanwhile=mlp_in)
            else:
                attn = torch.cat(attend(h, None, scales, None, rope_s, 0, _cabi.ATTN_BOTH, meanwhile=mlp_in)[:2], 1)
            g = box["g"]
            last = i + 1 == p.depth_single_blocks
            on = 0 if last else off(f"single_blocks.{i + 1}.modulation.lin")
            xm_next = None if last else torch.empty_like(x)

            def rows(r0, r1, o=o, on=on, last=last, xm_next=xm_next, attn=attn, g=g, pre=pre):
                # linear2 over cat(attn, gelu(mlp)) as two accumulating GEMMs (row-parallel when sharded)
                out = self._lin_rp(pre + ".linear2a", attn[:, r0:r1], bias_key=pre + ".linear2.bias",
                                   extra=(g[:, r0:r1], self.sd[pre + ".linear2b.weight"]))
                if last:
                    ops.gate_residual_ln_modulate(x[:, r0:r1], out, m_vec, None, None, o + 2 * D, 0, 0,
                                                  want_out=False)
                else:
                    ops.gate_residual_ln_modulate(x[:, r0:r1], out, m_vec, None, None, o + 2 * D, on, on + D,
                                                  out=xm_next[:, r0:r1])

            chunked(N, rows)
            xm = xm_next
        x = x[:, :P].contiguous()
        o = off("final_layer.adaLN_modulation.1")                                 # shift | scale
        pred = self._lin("final_layer.linear", ops.ln_modulate(x, m_vec, None, None, o, o + D))
        if seqpar:
            pred = self.hs.all_gather(pred, 1)               # [1,P_local,*] -> all rows, rank order
        return pred, collected

    __call__ = forward
