#!/usr/bin/env python
"""Benchmark of the ConceptAttention hot path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--config NAME]

metric  : concept-heatmap images/sec @1024^2 Flux-schnell (configs[1] of BASELINE.json:
          4096 image tokens, 256 text tokens, 4 concepts, 4 denoising steps, bf16,
          heatmaps from layers 15..18 of every step, both spaces).
step    : one image = 4 x (19 double + 38 single blocks) + heatmap accumulation.
value   : images/s with the step's inputs already resident in HBM (CUDA events), kernels launched
          one by one with a CUDA-event pair around every K2 launch (graph_replay: same, replayed).
e2e     : images/s through ConceptAttentionFluxPipeline.generate_image with pinned
          HOST inputs (latents, text/concept embeddings) copied H2D and the heatmaps
          copied D2H inside the timed region.
roofline: the fused joint-attention kernel (K2), timed live with CUDA events around
          every launch of the timed region, against the measured bf16 peak.
N > 1   : one process per GPU, each generates its own images (data parallel, no
          collective on the path); NCCL is used for the barrier / max-over-ranks only.
roofline_hbm : K1a / K1b / K1c / K3+K4, each launch of the timed region bracketed by CUDA events
          on its stream, algorithmic bytes / time against the measured copy bandwidth.
gpu_reference (N = 1): the reference's OWN modules (oracle/_ref) in bf16 on the same GPU, the full
          workload image through its denoise + compute_heatmaps_from_vectors — the real bar.
other_configs : the other BASELINE.json configurations measured in the same run (short).
--impl reference : the reference's own modules (oracle/_ref; the oracle port if absent) in fp32
          on all host cores, bounded sample of the workload.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

METRIC = "concept-heatmap images/sec @1024^2 Flux-schnell"
CONCEPT_WORDS = ["dragon", "rock", "sky", "sun", "tree", "grass", "cloud", "river", "bird", "stone",
                 "water", "fire", "snow", "sand", "road", "wall", "roof", "door", "lamp", "boat",
                 "fish", "leaf", "hill", "moon", "star", "rain", "wind", "path", "gate", "pond",
                 "reed", "moss"]
WORKLOADS = {
    # BASELINE.json configs[1] — the bench line.  mode generate: one step = one image.
    "schnell-1024": dict(model="flux-schnell", side=1024, C=4, nsteps=4, layers=list(range(15, 19)),
                         mode="generate", batch=1),
    # configs[2]: encode_image segmentation sweep, one noised timestep, 19 double blocks only;
    # one step = one batch of `batch` 512x512 images per GPU (data parallel).
    "encode-512": dict(model="flux-schnell", side=512, C=2, nsteps=4, layers=list(range(15, 19)),
                       mode="encode", batch=16),
    # configs[3]: flux-dev architecture, 20 concepts, 28 steps, per-layer AND per-timestep maps
    "dev-1024-c20": dict(model="flux-dev", side=1024, C=20, nsteps=28, layers=list(range(19)),
                         mode="generate", batch=1, per_pair=True, guidance=3.5),
    # configs[4]: 2048^2, 32 concepts, heads sharded over the ranks (strong scaling: all ranks
    # work on the same image; needs WORLD_SIZE dividing 24)
    "schnell-2048-hs": dict(model="flux-schnell", side=2048, C=32, nsteps=4, layers=list(range(15, 19)),
                            mode="generate", batch=1, head_shard=True),
    "tiny-256": dict(model="tiny", side=256, C=4, nsteps=2, layers=[1, 2], mode="generate", batch=1),
}
D_HEAD = 128


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return dict(tflops=p["bf16_tflops_sustained"], tflops_burst=p["bf16_tflops"],
                    hbm=p["hbm_gbs"], source="measured (MEASURED_PEAKS.json, sustained)")
    return dict(tflops=1400.0, tflops_burst=1590.0, hbm=6650.0, source="fallback (B200_PROFILING.md)")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "200", "-i", str(self.idx)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=lambda: self.lines.extend(self.proc.stdout), daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "samples": len(sm),
                "reasons": sorted(reasons)}


def attention_flops(P, T, C, H, double):
    """ALGORITHMIC FLOPs of one K2 launch per batch element."""
    f = 4 * D_HEAD * H * (T + P) ** 2
    return f + (4 * D_HEAD * H * C * (C + P) if double else 0)


def run_ours(args, rank, world, local_rank, secondary=False):
    from conceptattention_b200 import ConceptAttentionFluxPipeline, ops, sampling
    from conceptattention_b200.parallel import HeadShard, gather_results
    wl = WORKLOADS[args.config]
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    side, C, nsteps, layers = wl["side"], wl["C"], wl["nsteps"], wl["layers"]
    mode, batch = wl["mode"], wl["batch"]
    sharded = bool(wl.get("head_shard")) and world > 1
    if args.hs_mode == "auto":
        # sequence-parallel rows: the same GEMM rows and per-head attention as one GPU, computed
        # elsewhere (prediction bit-identical); the tensor-parallel form is ~4 % faster on 2 GPUs
        # but its row-parallel partial sums are a different bf16 evaluation (cross maps 2e-2 off)
        args.hs_mode = "sequence"
    guidance = wl.get("guidance", 0.0)
    concepts = CONCEPT_WORDS[:C]
    prompt = "A dragon standing on a rock"
    # head-sharded ranks hold the SAME weights and work on the same image; DP ranks differ
    hs = None
    if sharded:
        from conceptattention_b200.weights import CONFIGS
        hs = HeadShard(rank, world, CONFIGS[wl["model"]].num_heads,
                       sequence_parallel=args.hs_mode == "sequence",
                       peer_exchange=False if getattr(args, "no_peer_exchange", False) else None)
    from conceptattention_b200.weights import CONFIGS, init_state_dict
    # random-init weights with the reference's keys, drawn on the device; the same tensors are
    # handed to the reference modules of the gpu_reference arm (shared storage)
    sd = init_state_dict(CONFIGS[wl["model"]], seed=0 if sharded else rank, device=dev, dtype=torch.bfloat16)
    pipe = ConceptAttentionFluxPipeline(model_name=wl["model"], device=dev, state_dict=dict(sd), head_shard=hs)
    want_gpu_ref = (world == 1 and mode == "generate" and not args.no_gpu_reference and not secondary)
    if not want_gpu_ref:
        del sd
    gen = pipe.flux_generator
    pm = gen.params
    # text encoders on the HOST (pinned cache, pinned stacking buffers) so the e2e arm pays an
    # asynchronous H2D copy of their outputs
    host_t5 = sampling.SyntheticEmbedder(pm.context_in_dim, gen.t5.max_length, "cpu", pin=True)
    host_clip = sampling.SyntheticEmbedder(pm.vec_in_dim, None, "cpu", pin=True)
    for e, texts in ((host_t5, [prompt, ""] + concepts), (host_clip, [prompt, "", " ".join(concepts)])):
        for t in texts:
            e(t)
    seed0 = 0 if sharded else 1000 * rank
    host_lat = [torch.randn(batch, 16, side // 8, side // 8, generator=torch.Generator().manual_seed(seed0 + i))
                .to(torch.bfloat16).pin_memory() for i in range(args.steps + args.warmup)]
    dev_lat = [l.to(dev) for l in host_lat]
    P, T, H = (side // 16) ** 2, gen.t5.max_length, pm.num_heads
    sched = sampling.get_schedule(nsteps, P, shift=(not gen.is_schnell))

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    # ---- arm 1: inputs resident in HBM ------------------------------------------------
    if mode == "generate":
        inps = [gen.prepare_inputs(prompt, concepts, l) for l in dev_lat]

        def device_step(i):
            acc = gen.accumulator(1, C, P, layers, list(range(nsteps)), True, "sparsemax", False,
                                  per_pair=bool(wl.get("per_pair")))
            gen.denoise(inps[i], sched, guidance, acc)
            return acc.finalize()
    else:
        def device_step(i):
            out = pipe.encode_image(dev_lat[i], concepts, prompt="", width=side, height=side,
                                    layer_indices=layers, num_steps=nsteps, return_device_tensors=True)
            return out.concept_heatmaps, out.cross_attention_maps

    gen.model.use_graphs = not args.no_graphs
    if args.no_side_stream:
        gen.model.use_side_stream = False
    if args.tp_overlap:
        gen.model.overlap_allreduce = True
    for i in range(args.warmup):
        device_step(i)
    # graph capture empties the caching allocator's pools (torch.cuda.graph calls empty_cache):
    # one more untimed launch-by-launch step re-warms them for the timed region below
    ops.ATTN_EVENTS = []
    device_step(0)
    ops.ATTN_EVENTS = None
    barrier()
    # Timed region: K steps launch by launch with a CUDA-event pair around every K2 launch on
    # its stream, so `roofline` is measured inside the region `value` is measured on.  (While
    # ATTN_EVENTS is set the engine does not replay captured graphs: events recorded inside a
    # graph cannot be read back.  The graph-replay number is reported beside it.)
    ops.LAUNCHES = 0
    ops.ATTN_EVENTS = []
    clocks = ClockSampler(local_rank)
    clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.profiler.start()      # no-op unless run under `ncu --profile-from-start off`
    e0.record()
    for i in range(args.steps):
        device_step(args.warmup + i)
    e1.record()
    barrier()
    torch.cuda.profiler.stop()
    ms_dev = e0.elapsed_time(e1)
    launches = ops.LAUNCHES
    attn_ms = sum(ev[0].elapsed_time(ev[1]) for ev, *_ in ops.ATTN_EVENTS)
    n_attn = len(ops.ATTN_EVENTS)
    attn_flops = sum(B * attention_flops(Pq, Tq, Cq, Hq, Cq > 0) for _, B, Pq, Tq, Cq, Hq in ops.ATTN_EVENTS)
    ops.ATTN_EVENTS = None
    # the HBM-bound kernels, timed live the same way in a pass of their own (the same K steps again,
    # launch by launch, a CUDA-event pair around every K1a / K1b / K1c / K3+K4 launch): kept out of
    # the `value` region so that ~50 extra event records per block do not slow the number reported
    ops.HBM_EVENTS = {}
    e4, e5 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e4.record()
    for i in range(args.steps):
        device_step(args.warmup + i)
    e5.record()
    barrier()
    ms_hbm_pass = e4.elapsed_time(e5)
    pk = peaks()
    hbm = {}
    for name, evs in ops.HBM_EVENTS.items():
        ms_k = sum(ev[0].elapsed_time(ev[1]) for ev, _ in evs)
        by = sum(nb for _, nb in evs)
        gbs = by / (ms_k * 1e-3) / 1e9 if ms_k > 0 else 0.0
        hbm[name] = {"achieved": gbs, "frac": gbs / pk["hbm"], "launches_timed": len(evs),
                     "avg_launch_us": 1e3 * ms_k / max(len(evs), 1), "algorithmic_MB_per_launch": by / max(len(evs), 1) / 1e6,
                     "share_of_step": ms_k / ms_hbm_pass}
    ops.HBM_EVENTS = None
    # the same K steps as the product runs them by default (whole forwards replayed as CUDA graphs)
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record()
    for i in range(args.steps):
        device_step(args.warmup + i)
    e3.record()
    barrier()
    ms_graph = e2.elapsed_time(e3)
    clk = clocks.stop()
    graphed = gen.model.use_graphs and len(gen.model._graphs) > 0

    # ---- arm 2: end to end through the public API with host buffers ---------------------
    gen.t5, gen.clip = host_t5, host_clip

    def e2e_step(i):
        if mode == "generate":
            o = pipe.generate_image(prompt, concepts, width=side, height=side, layer_indices=layers,
                                    return_pil_heatmaps=False, num_inference_steps=nsteps,
                                    guidance=guidance, latents=host_lat[i], decode_image=False)
        else:
            o = pipe.encode_image(host_lat[i], concepts, prompt="", width=side, height=side,
                                  layer_indices=layers, num_steps=nsteps, return_pil_heatmaps=False)
        return o
    def gather_maps(outs_):
        # data parallel: the only communication of the run — gather the per-image maps on rank 0
        if world == 1 or sharded:
            return
        local = torch.as_tensor(np.stack([np.asarray(o.concept_heatmaps).reshape(batch, C, -1)
                                          for o in outs_])).reshape(len(outs_) * batch, C, -1).to(dev)
        gather_results(local, len(outs_) * batch * world, rank, world, dst=0)

    gather_maps([e2e_step(i) for i in range(args.warmup)])     # also creates the NCCL communicator
    barrier()
    t0 = time.perf_counter()
    e0.record()
    outs = []
    for i in range(args.steps):
        outs.append(e2e_step(args.warmup + i))
    out = outs[-1]
    gather_maps(outs)
    e1.record()
    barrier()
    ms_e2e = max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3)
    nprompt = 1 if mode == "generate" else batch
    h2d = (host_lat[0].numel() * 2 + nprompt * (T * pm.context_in_dim * 2 + pm.vec_in_dim * 2) + pm.vec_in_dim * 2
           + C * pm.context_in_dim * 2)
    d2h = 2 * batch * C * P * 4

    if world > 1:
        t = torch.tensor([ms_dev, ms_e2e, ms_graph], device=dev, dtype=torch.float64)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        ms_dev, ms_e2e, ms_graph = t.tolist()
    ach = attn_flops / (attn_ms * 1e-3) / 1e12 if attn_ms > 0 else 0.0
    # DRAM bytes per launch of K2 cannot be counted from inside the process: the figure is the one
    # `ncu --set full` capture of this command committed under profiles/ (file named in the line)
    traffic, traffic_src = None, None
    tpath = os.path.join(ROOT, "profiles", "k2_traffic.json")
    if os.path.exists(tpath) and args.config == "schnell-1024":
        with open(tpath) as f:
            tj = json.load(f)
        traffic, traffic_src = tj.get("dram_bytes_per_launch"), tj.get("source")
    images = args.steps * batch * (1 if sharded else world)
    blocks = "19 double + 38 single blocks" if mode == "generate" else "19 double blocks (stop_after_multimodal_attentions)"
    line = {
        "metric": METRIC, "value": images / (ms_dev * 1e-3), "unit": "images/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_dev / args.steps, "higher_is_better": True,
        "scaling": "strong" if sharded else "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic (random-init weights, randn latents/embeddings)",
        "config": {"workload": f"{args.config}: {wl['model']} {side}x{side}, P={P} T={T} C={C}, "
                               f"{mode}: {nsteps if mode == 'generate' else 1} DiT pass(es) x ({blocks}), heatmaps from "
                               f"layers {layers[0]}..{layers[-1]} both spaces",
                   "images_per_step_per_gpu": batch,
                   "parallelism": ((f"heads sharded {H}/{world} per GPU inside the attention, every token-wise op "
                                    f"on 1/{world} of the rows with replicated weights; two NCCL all-to-alls "
                                    f"per block") if args.hs_mode == "sequence" else
                                   (f"heads sharded {H}/{world} per GPU with the linears tensor-parallel around them; "
                                    f"NCCL all-reduce of partial heatmap logits and of the row-parallel "
                                    f"proj/MLP partial sums")) if sharded else
                                  f"dp{world} (one image stream per GPU, no collective on the data path)",
                   "cache": "inputs larger than L2: 23.8 GB of bf16 weights stream through every step"},
        "e2e": {"value": images / (ms_e2e * 1e-3), "unit": "images/s",
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e / args.steps},
        "gpu_launches": launches,
        "clocks": clk,
        "roofline": {"kernel": "ca_joint_attention (K2, tcgen05 one-way joint attention)",
                     "bound": "tensor", "achieved": ach, "peak": pk["tflops"], "unit": "TFLOP/s",
                     "frac": ach / pk["tflops"], "peak_source": pk["source"], "traffic": traffic,
                     "traffic_source": traffic_src, "frac_of_burst_peak": ach / pk["tflops_burst"],
                     "launches_timed": n_attn, "avg_launch_ms": attn_ms / max(n_attn, 1),
                     "share_of_step": attn_ms / ms_dev},
        "roofline_hbm": {"bound": "hbm", "peak": pk["hbm"], "unit": "GB/s", "peak_source": pk["source"],
                         "note": "a pass of the same K steps right after the timed region, every launch bracketed "
                                 "by CUDA events on its stream (rows = token rows per launch); "
                                 "algorithmic bytes per SURVEY.md 8d; at this size the operands are partly "
                                 "L2-resident (the producer kernel just wrote them), so frac can exceed what HBM "
                                 "alone would allow; the text stream's K1c launches overlap image-stream GEMMs",
                         "kernels": hbm},
        "graph_replay": {"enabled": graphed, "value": images / (ms_graph * 1e-3), "unit": "images/s",
                         "ms_per_step": ms_graph / args.steps,
                         "note": "same K steps, inputs resident, whole forwards replayed as CUDA graphs "
                                 "(the default of the public API, used by e2e)"},
    }
    if want_gpu_ref:
        ours_maps = (np.asarray(out.concept_heatmaps), np.asarray(out.cross_attention_maps))
        gen.model._graphs.clear()                       # free the captured graphs' private pools
        torch.cuda.empty_cache()
        # the last timed e2e step used host_lat[warmup + steps - 1] == dev_lat[-1]
        gr = run_gpu_reference(wl, sd, gen, dev, dev_lat, prompt, concepts, ours_maps)
        gr["speedup_e2e_over_gpu_reference"] = line["e2e"]["value"] / gr["value"]
        gr["speedup_value_over_gpu_reference"] = line["value"] / gr["value"]
        line["gpu_reference"] = gr
    return line, out


def run_reference(args):
    """CPU arm: the reference's OWN modules (oracle/_ref: ModifiedDoubleStreamBlock,
    ModifiedSingleStreamBlock, compute_heatmaps_from_vectors — unmodified files; the oracle port
    only if oracle/_ref is absent) in fp32 on all host cores, default nn.Module init, synthetic
    inputs of the workload's shapes.  One step = a bounded SAMPLE of one image of the workload:
    1 double block (incl. the full (C+P)^2 second attention it executes) + 1 single block + the
    heatmap computation of one (timestep, 4 layers) slice at full dimensions, extrapolated linearly:
        t_image = nsteps * (depth t_double + depth_single t_single) + nsteps * t_heat
    (the model-independent glue — embeddings, final layer — is < 0.1 %)."""
    import oracle
    from oracle import reference_modules as rm
    from conceptattention_b200.weights import CONFIGS, FluxParams, init_state_dict
    wl = WORKLOADS[args.config]
    pm = CONFIGS[wl["model"]]
    cores = os.cpu_count()
    torch.set_num_threads(cores)
    torch.set_grad_enabled(False)
    side, C, nsteps, layers = wl["side"], wl["C"], wl["nsteps"], wl["layers"]
    D, H = pm.hidden_size, pm.num_heads
    P, T = (side // 16) ** 2, {"flux-schnell": 256, "flux-dev": 512}.get(wl["model"], 128)
    g = torch.Generator().manual_seed(0)
    img, txt, con = (torch.randn(1, n, D, generator=g) for n in (P, T, C))
    vec, cvec = torch.randn(1, D, generator=g), torch.randn(1, D, generator=g)
    ids = oracle.image_ids(1, side // 16, side // 16)
    nl = min(len(layers), 4)
    use_ref = rm.available()
    if use_ref:
        ns = rm.load()
        from concept_attention.flux.src.flux.modules.layers import EmbedND
        torch.manual_seed(0)
        dbl = ns.ModifiedDoubleStreamBlock(D, H, mlp_ratio=pm.mlp_ratio, qkv_bias=pm.qkv_bias).eval()
        sgl = ns.ModifiedSingleStreamBlock(D, H, mlp_ratio=pm.mlp_ratio).eval()
        embed = EmbedND(dim=D // H, theta=pm.theta, axes_dim=pm.axes_dim)
        pe = embed(torch.cat((torch.zeros(1, T, 3), ids), 1))
        pe_c = embed(torch.cat((torch.zeros(1, C, 3), ids), 1))
        kind, what = "reference", f"the reference's own modules ({ns.kind})"
    else:
        one = FluxParams(**{**pm.asdict(), "depth": 1, "depth_single_blocks": 1})
        sd = init_state_dict(one, seed=0)
        pe = oracle.rope_table(torch.cat((torch.zeros(1, T, 3), ids), 1), pm.axes_dim, pm.theta)
        pe_c = oracle.rope_table(torch.cat((torch.zeros(1, C, 3), ids), 1), pm.axes_dim, pm.theta)
        kind, what = "port", "the oracle port (oracle/_ref absent)"

    def sample():
        t0 = time.perf_counter()
        if use_ref:
            i2, t2, c2, vecs = dbl(img=img, txt=txt, vec=vec, pe=pe, concepts=con, concept_vec=cvec, concept_pe=pe_c)
        else:
            i2, t2, c2, vecs = oracle.double_block(sd, "double_blocks.0", H, img, txt, vec, pe, con, cvec, pe_c)
        t1 = time.perf_counter()
        if use_ref:
            sgl(torch.cat((t2, i2), 1), vec=vec, pe=pe)
        else:
            oracle.single_block(sd, "single_blocks.0", H, torch.cat((t2, i2), 1), vec, pe)
        t2_ = time.perf_counter()
        for a, b in (("output_space_image_vectors", "output_space_concept_vectors"),
                     ("cross_attention_image_vectors", "cross_attention_concept_vectors")):
            iv = vecs[a][None, None].expand(1, nl, *vecs[a].shape)
            cv = vecs[b][None, None].expand(1, nl, *vecs[b].shape)
            if use_ref:
                ns.compute_heatmaps_from_vectors(iv, cv, layer_indices=list(range(nl)), timesteps=[0])
            else:
                oracle.heatmaps_from_vectors(iv, cv, list(range(nl)), [0])
        t3 = time.perf_counter()
        if wl["mode"] == "encode":            # one pass through the 19 double blocks + heatmaps
            return pm.depth * (t1 - t0) + (t3 - t2_)
        return nsteps * (pm.depth * (t1 - t0) + pm.depth_single_blocks * (t2_ - t1)) + nsteps * (t3 - t2_)

    w0 = time.perf_counter()
    for _ in range(max(args.warmup, 1)):
        sample()
    per_sample = (time.perf_counter() - w0) / max(args.warmup, 1)
    steps = max(1, min(args.steps, int(150.0 / max(per_sample, 1e-3))))
    est = [sample() for _ in range(steps)]
    t_img = statistics.mean(est)
    val = 1.0 / t_img
    desc = (f"{what}, fp32, {cores} threads; per step: 1 double block + 1 single block + heatmaps of 1 timestep x "
            f"{nl} layers x 2 spaces at full size (P={P},T={T},C={C},D={D}), extrapolated linearly to {nsteps} steps x "
            f"({pm.depth}+{pm.depth_single_blocks}) blocks (not a whole image: a whole image is ~{t_img:.0f} s of CPU time)")
    return {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "images/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": max(args.warmup, 1), "ms_per_step": t_img * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic (random-init weights, randn inputs)",
        "config": {"workload": f"{args.config}: {wl['model']} {side}x{side}, P={P} T={T} C={C}, {nsteps} denoise steps",
                   "note": "CPU host cores; rank 0 only; bounded sample extrapolated (ms_per_step is the "
                           "extrapolated time of one image, not the wall time of a step)"},
        "cpu_baseline": {"value": val, "unit": "images/s", "cores": cores, "kind": kind, "sample": desc,
                         "sample_wall_s": per_sample},
        "e2e": {"value": val, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }


def run_gpu_reference(wl, sd, gen, dev, dev_lat, prompt, concepts, ours_maps, n_images=2):
    """The real bar: the reference's own PyTorch path in bf16 on the
    SAME B200, same random-init weights (shared storage), same inputs: its `denoise`
    (flux/src/flux/sampling.py:96-152: ModifiedFluxDiT with two full SDPA calls per double block,
    the per-layer / per-step torch.stack's) followed by `compute_heatmaps_from_vectors` for both
    spaces and the fp32 host copy of the maps (concept_attention_pipeline.py:150-172).  Inputs are
    resident in HBM, like `value`.  Not timed: T5/CLIP (synthetic in both arms), the VAE, and the
    reference generator's model.cpu()/.to(device) bounce per image (image_generator.py:183-194),
    which would only make it slower."""
    import oracle
    from oracle import reference_modules as rm
    from conceptattention_b200 import sampling
    pm = gen.params
    side, nsteps, layers = wl["side"], wl["nsteps"], wl["layers"]
    P = (side // 16) ** 2
    sched = sampling.get_schedule(nsteps, P, shift=(not gen.is_schnell))
    guidance = wl.get("guidance", 0.0)
    if rm.available():
        ns = rm.load()
        model = rm.build_dit(pm.asdict(), sd, dev, torch.bfloat16)
        kind = f"the reference's own modules ({ns.kind}), bf16, cuda"

        def image(lat):
            inp = gen.prepare_inputs(prompt, concepts, lat)
            x, _, st = ns.denoise(model, img=inp["img"], img_ids=inp["img_ids"], txt=inp["txt"],
                                  txt_ids=inp["txt_ids"], vec=inp["vec"], timesteps=sched, guidance=guidance,
                                  concepts=inp["concepts"], concept_ids=inp["concept_ids"],
                                  concept_vec=inp["concept_vec"])
            out, cross = rm.heatmaps(st, layers, range(nsteps))
            return out.to(torch.float32).cpu().numpy(), cross.to(torch.float32).cpu().numpy()
    else:
        sdd = {k: v.to(dev, torch.bfloat16) for k, v in sd.items()}
        kind = "the oracle port of the reference op sequence (oracle/_ref absent), bf16, cuda"

        def image(lat):
            inp = gen.prepare_inputs(prompt, concepts, lat)
            x, st = oracle.denoise(sdd, pm.asdict(), inp["img"], inp["img_ids"].float(), inp["txt"],
                                   inp["txt_ids"].float(), inp["vec"], sched, guidance=guidance,
                                   concepts=inp["concepts"], concept_ids=inp["concept_ids"].float(),
                                   concept_vec=inp["concept_vec"])
            maps = [oracle.heatmaps_from_vectors(st[a], st[b], layers, range(nsteps)) for a, b in
                    (("output_space_image_vectors", "output_space_concept_vectors"),
                     ("cross_attention_image_vectors", "cross_attention_concept_vectors"))]
            return maps[0].float().cpu().numpy(), maps[1].float().cpu().numpy()

    with torch.no_grad():
        image(dev_lat[0])                       # warm-up (cuBLAS heuristics, SDPA backend selection, allocator)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(n_images):                # the LAST timed image is dev_lat[-1] (parity below)
            maps = image(dev_lat[-((n_images - 1 - i) % len(dev_lat)) - 1])
        e1.record()
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n_images
    # parity of the two arms on the last image (same latents as the last timed step of `value`)
    par = {}
    if ours_maps is not None:
        for tag, got, ref in (("concept_heatmaps", ours_maps[0], maps[0]), ("cross_attention_maps", ours_maps[1], maps[1])):
            got, ref = np.asarray(got, dtype=np.float32).reshape(ref.shape), np.asarray(ref, dtype=np.float32)
            par[tag] = {"max_rel": float((np.abs(got - ref) / np.maximum(np.abs(ref), 1e-2)).max()),
                        "mean_abs": float(np.abs(got - ref).mean()),
                        "argmax_mask_agreement": float((got.argmax(-3) == ref.argmax(-3)).mean())}
    return {"value": 1e3 / ms, "unit": "images/s", "ms_per_image": ms, "images_timed": n_images, "kind": kind,
            "what": "denoise (all DiT blocks, two SDPA calls per double block, vector stacks) + "
                    "compute_heatmaps_from_vectors x 2 spaces + fp32 host copy; inputs resident in HBM",
            "parity_vs_ours_same_image": par}


def other_configs(args, rank, world, local_rank):
    """Short runs of the other BASELINE.json configurations in the same process (extra keys of
    the line; the bench line itself stays configs[1]).  N = 1: configs[2] (encode sweep),
    configs[3] (flux-dev, 20 concepts, 28 steps, per-(timestep, layer) maps) and configs[4] on one GPU
    (the strong-scaling base).  N > 1: configs[2] data-parallel and configs[4] with the heads
    sharded over the N ranks (the curve that has a collective in it)."""
    import gc
    out = {}
    todo = [("encode-512", 3, 3), ("schnell-2048-hs", 2, 3)]
    if world == 1:
        todo.insert(1, ("dev-1024-c20", 1, 3))
    for name, steps, warmup in todo:
        if name == "schnell-2048-hs" and world > 1 and 24 % world:
            continue
        gc.collect()
        torch.cuda.empty_cache()
        a = argparse.Namespace(**{**vars(args), "config": name, "steps": steps, "warmup": warmup,
                                  "hs_mode": "auto"})
        try:
            ln, _ = run_ours(a, rank, world, local_rank, secondary=True)
        except Exception as e:                              # report, do not lose the bench line
            out[name] = {"error": f"{type(e).__name__}: {e}"[:300]}
            continue
        out[name] = {k: ln[k] for k in ("value", "unit", "ms_per_step", "steps", "scaling", "n_gpus")}
        out[name].update(workload=ln["config"]["workload"], parallelism=ln["config"]["parallelism"],
                         images_per_step_per_gpu=ln["config"]["images_per_step_per_gpu"],
                         e2e=ln["e2e"]["value"], graph_replay=ln["graph_replay"]["value"],
                         k2_tflops=ln["roofline"]["achieved"], k2_frac=ln["roofline"]["frac"],
                         k2_avg_launch_ms=ln["roofline"]["avg_launch_ms"],
                         hbm_kernels={k: round(v["frac"], 3) for k, v in ln["roofline_hbm"]["kernels"].items()},
                         clocks=ln["clocks"])
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="schnell-1024", choices=list(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graphs", action="store_true", help="launch every kernel individually (A/B)")
    ap.add_argument("--hs-mode", choices=("auto", "tensor", "sequence"), default="auto",
                    help="head-sharded workloads: tensor-parallel linears, or sequence-parallel rows "
                         "with the head split only inside the attention; auto = sequence on more than "
                         "2 GPUs (measured: 4.22 vs 3.19 images/s on 8, 1.22 vs 1.27 on 2)")
    ap.add_argument("--tp-overlap", action="store_true",
                    help="head-sharded mode: two token chunks on two streams around the all-reduces (A/B)")
    ap.add_argument("--no-peer-exchange", action="store_true",
                    help="sequence-parallel exchanges through NCCL all-to-all instead of peer memory (A/B)")
    ap.add_argument("--no-side-stream", action="store_true", help="text stream on the main CUDA stream (A/B)")
    ap.add_argument("--no-gpu-reference", action="store_true", help="skip the reference-on-the-same-GPU arm")
    ap.add_argument("--no-other-configs", action="store_true", help="skip the short runs of the other configs")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        if rank == 0:
            print(json.dumps(run_reference(args)), flush=True)
        return
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    args.warmup = max(args.warmup, 3)
    line, _ = run_ours(args, rank, world, local_rank)
    if args.config == "schnell-1024" and not args.no_other_configs:
        line["other_configs"] = other_configs(args, rank, world, local_rank)
    if rank == 0:
        if world == 1 and not args.no_cpu_baseline:
            torch.cuda.empty_cache()
            ref = run_reference(argparse.Namespace(config=args.config, steps=2, warmup=1, gpus=1))
            line["cpu_baseline"] = ref["cpu_baseline"]
        print(json.dumps(line), flush=True)
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
