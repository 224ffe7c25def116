"""ConceptAttentionFluxPipeline — drop-in API of the reference pipeline
(concept_attention/concept_attention_pipeline.py:20-24,93-357) over the B200
engine.  Signatures, defaults and error behaviour of `generate_image` /
`encode_image` / `compute_heatmaps_from_vectors` are kept; what changes is where
the work happens: vectors never leave the device, heatmaps are accumulated block
by block (HeatmapAccumulator), no debug sync, no model.cpu() round trip.

Offline there are no Flux / T5 / CLIP / VAE checkpoints: weights are random-init
with the reference's state-dict keys (weights.init_state_dict) unless a state
dict is supplied, and the text encoders are `SyntheticEmbedder`s.
"""
from __future__ import annotations

import math
import os
import time
from dataclasses import dataclass

import numpy as np
import torch

from . import _cabi, ops, sampling
from .dit import FluxDiTEngine, HeatmapAccumulator
from .weights import CONFIGS, T5_MAX_LENGTH, init_state_dict, load_checkpoint

try:                                     # PIL is optional at import time
    import PIL.Image
except ImportError:                      # pragma: no cover
    PIL = None


@dataclass
class ConceptAttentionPipelineOutput:
    image: "PIL.Image.Image | np.ndarray"
    concept_heatmaps: "list[PIL.Image.Image] | np.ndarray"
    cross_attention_maps: "list[PIL.Image.Image] | np.ndarray"
    # extension (return_per_pair=True): fp32 device tensors [Nt_sel, L_sel, B, C, h, w] — every
    # selected (timestep, layer) pair's map before the mean (concept_attention_pipeline.py:76-77)
    per_pair_concept_heatmaps: "torch.Tensor | None" = None
    per_pair_cross_attention_maps: "torch.Tensor | None" = None


def compute_heatmaps_from_vectors(image_vectors, concept_vectors, layer_indices,
                                  timesteps=list(range(4)), softmax=True, normalize_concepts=False,
                                  attention_norm="sparsemax", round_bf16=False):
    """Same call as the reference function (concept_attention_pipeline.py:29-91) on
    stacked vectors [Nt,L,B,P,D] / [Nt,L,B,H,P,d] (bf16, CUDA).  Returns [B,C,h,w]
    fp32 with h = w = isqrt(P) (the reference hard-codes 64)."""
    if not (softmax or attention_norm in ("softmax", "entmax15", "sparsemax")):
        raise ValueError(f"Unknown attention_norm={attention_norm}")
    norm = "softmax" if (softmax or attention_norm == "softmax") else attention_norm
    per_head = image_vectors.ndim == 6
    B, P, C = image_vectors.shape[2], image_vectors.shape[-2], concept_vectors.shape[-2]
    acc = torch.zeros(B, C, P, device=image_vectors.device, dtype=torch.float32)
    n = 0
    for t in timesteps:
        for l in layer_indices:
            iv, cv = image_vectors[int(t), l], concept_vectors[int(t), l]
            if normalize_concepts:
                # linear_normalization over the concept axis (utils.py:35-44), per batch
                if per_head:
                    cv = cv.permute(0, 2, 1, 3).reshape(B, C, -1)
                    iv = iv.permute(0, 2, 1, 3).reshape(B, P, -1)
                cv = torch.stack([ops.concept_linear_normalize(cv[b].contiguous()) for b in range(B)])
                ops.heatmap_accumulate(iv.contiguous(), cv, acc, norm, round_bf16)
            else:
                ops.heatmap_accumulate(iv, cv, acc, norm, round_bf16, per_head=per_head)
            n += 1
    return ops.finalize_heatmaps(acc, n, round_bf16)


class FluxGenerator:
    """Holds the engine, the (synthetic) text encoders and the optional VAE; mirrors
    concept_attention/image_generator.py:64-205."""

    def __init__(self, model_name="flux-schnell", device="cuda:0", offload=False,
                 state_dict=None, seed=0, t5=None, clip=None, ae=None, head_shard=None):
        self.device = torch.device(device)
        self.offload = offload
        self.model_name = model_name
        self.is_schnell = "schnell" in model_name or model_name == "tiny"
        self.params = CONFIGS[model_name]
        if isinstance(state_dict, (str, os.PathLike)):       # a flux1-*.safetensors / .pt file
            state_dict = load_checkpoint(state_dict, self.params, device=self.device)
        if state_dict is None:
            state_dict = init_state_dict(self.params, seed=seed, device=self.device,
                                         dtype=torch.bfloat16)
        self.model = FluxDiTEngine(self.params, state_dict, device=self.device, head_shard=head_shard)
        self.t5 = t5 or sampling.SyntheticEmbedder(self.params.context_in_dim,
                                                   T5_MAX_LENGTH[model_name], self.device)
        self.clip = clip or sampling.SyntheticEmbedder(self.params.vec_in_dim, None, self.device)
        self.ae = ae
        self._acc_buffers = {}
        self._id_cache = {}

    def accumulator(self, B, C, P, layer_indices, timesteps, softmax, attention_norm, round_bf16,
                    per_pair=False):
        """HeatmapAccumulator on generator-owned state buffers: the same addresses call after
        call, so the engine's captured forwards stay valid from image to image.  `per_pair`
        adds the [Nt_sel, L_sel, 2, B, C, P] fp32 per-(timestep, layer) buffer (config 4:
        28 x 19 x 2 x 20 x 4096 x 4 B = 349 MB, where the reference stacks 26.8 GB of vectors,
        flux/src/flux/sampling.py:146-150)."""
        def owned(key, shape, dtype=torch.float32):
            buf = self._acc_buffers.get(key)
            if buf is None:
                if len(self._acc_buffers) >= 8:
                    self._acc_buffers.pop(next(iter(self._acc_buffers)))
                buf = self._acc_buffers[key] = torch.empty(shape, device=self.device, dtype=dtype)
            return buf
        pairs = slot = None
        if per_pair:
            nt = len(set(int(t) for t in timesteps)) if timesteps is not None else 1
            nl = len(set(int(l) for l in layer_indices))
            pairs = owned(("pairs", nt, nl, B, C, P), (nt, nl, 2, B, C, P))
            slot = owned(("pair_slot",), (1,), torch.int32)
        return HeatmapAccumulator(B, C, P, self.device, layer_indices, timesteps, softmax,
                                  attention_norm, round_bf16, per_pair=per_pair,
                                  buffer=owned((B, C, P), (2, B, C, P)), pair_buffer=pairs, pair_slot=slot)

    @torch.inference_mode()
    def prepare_inputs(self, prompt, concepts, latents):
        inp = sampling.prepare(self.t5, self.clip, latents, prompt)
        emb, ids, vec = sampling.embed_concepts(self.clip, self.t5, concepts)
        B = inp["img"].shape[0]
        inp["concepts"] = emb.to(self.device, non_blocking=True).expand(B, -1, -1)
        inp["concept_ids"] = ids.to(self.device).expand(B, -1, -1)
        inp["concept_vec"] = vec.to(self.device, non_blocking=True).expand(B, -1)
        # position ids depend on the shapes only (sampling.py:40-47, utils.py:25): keep ONE set
        # per shape so the engine's RoPE-table cache hits by tensor identity (no D2H, no hash)
        key = tuple(tuple(inp[k].shape) for k in ("img_ids", "txt_ids", "concept_ids"))
        ids = self._id_cache.get(key)
        if ids is None:
            if len(self._id_cache) >= 8:
                self._id_cache.pop(next(iter(self._id_cache)))
            ids = self._id_cache[key] = tuple(inp[k].contiguous() for k in ("img_ids", "txt_ids", "concept_ids"))
        inp["img_ids"], inp["txt_ids"], inp["concept_ids"] = ids
        return inp

    @torch.inference_mode()
    def denoise(self, inp, timesteps, guidance, accumulator, joint_attention_kwargs=None):
        """Euler loop of flux/src/flux/sampling.py:96-152 feeding the accumulator
        instead of stacking vectors."""
        img = inp["img"]
        g = torch.full((img.shape[0],), guidance, device=self.device, dtype=img.dtype)
        for step, (t_curr, t_prev) in enumerate(zip(timesteps[:-1], timesteps[1:])):
            if accumulator is not None:
                accumulator.begin_step(step)
            t_vec = torch.full((img.shape[0],), t_curr, dtype=img.dtype, device=self.device)
            pred, _ = self.model(img=img, img_ids=inp["img_ids"], txt=inp["txt"],
                                 txt_ids=inp["txt_ids"], concepts=inp["concepts"],
                                 concept_ids=inp["concept_ids"], concept_vec=inp["concept_vec"],
                                 y=inp["vec"], timesteps=t_vec, guidance=g, iteration=step,
                                 joint_attention_kwargs=joint_attention_kwargs,
                                 accumulator=accumulator, collect_vectors=False)
            img = img + (t_prev - t_curr) * pred
        return img

    def decode(self, x, height, width):
        """Latents -> image.  With a VAE: ae.decode (image_generator.py:183-204, minus
        the model.cpu() bounce).  Without one (offline): returns the unpacked
        latents [16,h/8,w/8] as float32 numpy — documented stand-in."""
        lat = sampling.unpack(x.float(), height, width)
        if self.ae is None:
            return lat[0].cpu().numpy()
        with torch.autocast(device_type=self.device.type, dtype=torch.bfloat16):
            img = self.ae.decode(lat)
        img = img.clamp(-1, 1)[0].permute(1, 2, 0)
        arr = (127.5 * (img + 1.0)).cpu().byte().numpy()
        return PIL.Image.fromarray(arr) if PIL is not None else arr


def _colour(maps: np.ndarray, cmap: str):
    """Global min-max over all concepts, colormap, uint8 RGB, PIL
    (concept_attention_pipeline.py:174-196)."""
    lo, hi = maps.min(), maps.max()
    out = []
    for m in maps:
        rgb = sampling.apply_colormap((m - lo) / (hi - lo), cmap)
        out.append(PIL.Image.fromarray(rgb) if PIL is not None else rgb)
    return out


def _check_concepts(concepts):
    """The kernels hold one concept per register slot (CA_MAX_CONCEPTS = 32); the reference
    accepts any count.  Reject early with a clear message instead of a kernel status deep inside."""
    if not 1 <= len(concepts) <= _cabi.MAX_CONCEPTS:
        raise ValueError(f"between 1 and {_cabi.MAX_CONCEPTS} concepts are supported per call, got "
                         f"{len(concepts)}; split the concept list over several calls")


class ConceptAttentionFluxPipeline:
    """Generate images with Flux, or 'encode' existing ones, and return per-concept
    heatmaps.  Same constructor and methods as the reference class
    (concept_attention_pipeline.py:93-357); extra keyword arguments are optional
    and default to the reference behaviour."""

    def __init__(self, model_name: str = "flux-schnell", offload_model=False, device="cuda:0",
                 state_dict=None, seed=0, t5=None, clip=None, ae=None, head_shard=None):
        self.model_name = model_name
        self.offload_model = offload_model
        _cabi.load()                      # fail loudly before any weights are built
        self.flux_generator = FluxGenerator(model_name=model_name, offload=offload_model,
                                            device=device, state_dict=state_dict, seed=seed,
                                            t5=t5, clip=clip, ae=ae, head_shard=head_shard)

    # ------------------------------------------------------------------------------
    @torch.no_grad()
    def generate_image(self, prompt: str, concepts: list, width: int = 1024, height: int = 1024,
                       return_cross_attention=False, layer_indices=list(range(15, 19)),
                       return_pil_heatmaps=True, seed: int = 0, num_inference_steps: int = 4,
                       guidance: float = 0.0, timesteps=None, softmax: bool = True,
                       attention_norm: str = "sparsemax", cmap="plasma",
                       joint_attention_kwargs=None, latents=None, round_bf16=False,
                       decode_image=True, return_per_pair=False) -> ConceptAttentionPipelineOutput:
        gen = self.flux_generator
        depth = gen.params.depth
        assert return_cross_attention is False, "Not supported yet"
        assert all(0 <= i < depth for i in layer_indices), "Invalid layer index"
        assert height == width, "Height and width must be the same for now"
        _check_concepts(concepts)
        if timesteps is None:
            timesteps = list(range(num_inference_steps))
        dev = gen.device
        if latents is None:
            latents = sampling.get_noise(1, height, width, dev, torch.bfloat16, seed)
        else:                              # host latents: the only H2D copy of a generate call
            latents = latents.to(dev, torch.bfloat16, non_blocking=True)
        inp = gen.prepare_inputs(prompt, concepts, latents)
        sched = sampling.get_schedule(num_inference_steps, inp["img"].shape[1],
                                      shift=(not gen.is_schnell))
        B, P, C = inp["img"].shape[0], inp["img"].shape[1], len(concepts)
        acc = gen.accumulator(B, C, P, layer_indices, timesteps, softmax, attention_norm, round_bf16,
                              per_pair=return_per_pair)
        x = gen.denoise(inp, sched, guidance, acc, joint_attention_kwargs)
        concept_heatmaps, cross_attention_maps = acc.finalize()
        pp = acc.per_pair_heatmaps() if return_per_pair else (None, None)
        image = gen.decode(x, height, width) if decode_image else None
        # one D2H of [2,C,h,w] fp32 (the reference moves both maps separately, :171-172)
        both = torch.stack((concept_heatmaps[0], cross_attention_maps[0])).cpu().numpy()
        concept_heatmaps, cross_attention_maps = both[0], both[1]
        if return_pil_heatmaps:
            concept_heatmaps = _colour(concept_heatmaps, cmap)
            cross_attention_maps = _colour(cross_attention_maps, cmap)
        return ConceptAttentionPipelineOutput(image=image, concept_heatmaps=concept_heatmaps,
                                              cross_attention_maps=cross_attention_maps,
                                              per_pair_concept_heatmaps=pp[0],
                                              per_pair_cross_attention_maps=pp[1])

    # ------------------------------------------------------------------------------
    @torch.no_grad()
    def encode_image(self, image, concepts: list, prompt: str = "", width: int = 1024,
                     height: int = 1024, layer_indices=list(range(15, 19)), num_samples: int = 1,
                     num_steps: int = 4, noise_timestep: int = 2, device: str = "cuda:0",
                     return_pil_heatmaps: bool = True, seed: int = 0, cmap="plasma",
                     stop_after_multi_modal_attentions=True, attention_norm: str = "sparsemax",
                     softmax=True, joint_attention_kwargs=None, round_bf16=False,
                     return_device_tensors=False) -> ConceptAttentionPipelineOutput:
        """`image` is a PIL image (needs a VAE: `ae=` at construction) or a latent
        tensor [B,16,height/8,width/8] (the offline / benchmark path; B > 1 runs the
        whole batch through one forward).  Differences from the reference, all
        reference defects: width/height are honoured; every one of
        the `num_samples` noise draws is averaged (the reference indexes the sample
        axis with the float schedule tail, which truncates to sample 0)."""
        gen = self.flux_generator
        depth = gen.params.depth
        assert all(0 <= i < depth for i in layer_indices), "Invalid layer index"
        assert height == width, "Height and width must be the same for now"
        _check_concepts(concepts)
        dev = gen.device
        if isinstance(image, torch.Tensor):
            clean = image.to(dev, torch.bfloat16, non_blocking=True)
        else:
            if gen.ae is None:
                raise RuntimeError("encode_image(PIL image) needs a VAE (ae=...); offline pass "
                                   "latents [B,16,h/8,w/8] instead")
            # uint8 HWC goes to the device as it is; scale, resize (nearest, as torch's
            # F.interpolate default in segmentation.py:115-131) and VAE encode run there
            arr = torch.from_numpy(np.array(image.convert("RGB"))).to(dev, non_blocking=True)
            arr = arr.permute(2, 0, 1).float() / 255.0
            arr = torch.nn.functional.interpolate((2.0 * arr - 1.0)[None], (height, width))
            with torch.autocast(device_type=dev.type, dtype=torch.bfloat16):
                clean = gen.ae.encode(arr)
        B = clean.shape[0]
        acc = None
        for i in range(num_samples):
            noise = sampling.get_noise(1, height, width, dev, torch.bfloat16, seed + i)
            sched = sampling.get_schedule(num_steps, noise.shape[-1] * noise.shape[-2] // 4,
                                          shift=(not gen.is_schnell))
            t = sched[noise_timestep]
            noised = t * noise + (1.0 - t) * clean
            inp = gen.prepare_inputs(prompt if B == 1 else [prompt] * B, concepts, noised)
            if acc is None:
                acc = gen.accumulator(B, len(concepts), inp["img"].shape[1], layer_indices,
                                      None, softmax, attention_norm, round_bf16)
            t_vec = torch.full((B,), t, dtype=torch.bfloat16, device=dev)
            g_vec = torch.zeros(B, dtype=torch.bfloat16, device=dev)
            gen.model(img=inp["img"], img_ids=inp["img_ids"], txt=inp["txt"], txt_ids=inp["txt_ids"],
                      concepts=inp["concepts"], concept_ids=inp["concept_ids"],
                      concept_vec=inp["concept_vec"], y=inp["concept_vec"], timesteps=t_vec,
                      guidance=g_vec,
                      stop_after_multimodal_attentions=stop_after_multi_modal_attentions,
                      joint_attention_kwargs=joint_attention_kwargs, accumulator=acc,
                      collect_vectors=False)
        concept_heatmaps, cross_attention_maps = acc.finalize()
        if return_device_tensors:
            return ConceptAttentionPipelineOutput(image=image, concept_heatmaps=concept_heatmaps,
                                                  cross_attention_maps=cross_attention_maps)
        both = torch.stack((concept_heatmaps, cross_attention_maps)).cpu().numpy()
        if B == 1:
            both = both[:, 0]
        concept_heatmaps, cross_attention_maps = both[0], both[1]
        if return_pil_heatmaps and B == 1:
            concept_heatmaps = _colour(concept_heatmaps, cmap)
            cross_attention_maps = _colour(cross_attention_maps, cmap)
        return ConceptAttentionPipelineOutput(image=image, concept_heatmaps=concept_heatmaps,
                                              cross_attention_maps=cross_attention_maps)

    # ------------------------------------------------------------------------------
    @staticmethod
    def segment(concept_heatmaps, mode="argmax", num_background=0, apply_blur=False):
        """Heatmap -> mask on the device.  mode 'argmax': int64 [B,h,w]
        (multi_class_segmentation.py:71-77); mode 'mean': bool [B,C,h,w], coeff > mean
        (segmentation.py:56-61)."""
        heat = concept_heatmaps
        if not isinstance(heat, torch.Tensor):
            heat = torch.as_tensor(np.asarray(heat), dtype=torch.float32, device="cuda")
        if heat.ndim == 3:
            heat = heat[None]
        heat = heat.float().contiguous()
        if mode == "argmax":
            return ops.argmax_mask(heat, num_background, apply_blur)
        if mode == "mean":
            return ops.binary_masks(heat, apply_blur)
        raise ValueError(f"unknown mode {mode}")
