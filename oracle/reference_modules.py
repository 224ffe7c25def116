"""Loader for the reference's OWN modules (oracle/_ref, else /root/reference).
TEST / BASELINE INFRASTRUCTURE — see oracle/__init__.py; never imported by the product.

The reference's files are imported unmodified.  What this module adds around them:

* stubs for four third-party packages the reference imports eagerly and this image
  lacks, none on the hot path: `entmax` (raises if called: sparsemax/entmax15 stay
  "parity unpinned"), `matplotlib`, `imwatermark`, `fire`;
* three documented patches of reference defects, applied by
  shadowing module globals, not by editing files:
    (1) `compute_heatmaps_from_vectors` hard-codes h=w=64
        (concept_attention_pipeline.py:85-90): its `einops.rearrange` call gets
        h=w=isqrt(P), so sizes other than 1024x1024 work;
    (2) the debug `print` of concept_attention_pipeline.py:72-74 is silenced (the
        `.item()` sync it makes stays: it is reference behaviour);
    (3) `tqdm` progress bars of flux/src/flux/sampling.py:121 are silenced;
* `build_dit`: the reference `ModifiedFluxDiT` constructed on the meta device and filled
  from a state dict with `assign=True` — the reference's own loading path
  (image_generator.py:38-45) — so a full-size model costs no host RAM.
"""
from __future__ import annotations

import math
import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
_NS = None


def root():
    """Directory holding the `concept_attention` package, or None."""
    for r in (os.path.join(HERE, "_ref"), "/root/reference"):
        if os.path.exists(os.path.join(r, "concept_attention", "modified_flux_dit.py")):
            return r
    return None


def available() -> bool:
    return root() is not None


def install_stubs():
    def mod(name, **attrs):
        if name in sys.modules and not getattr(sys.modules[name], "_ca_stub", False):
            return sys.modules[name]              # the real package is installed: use it
        m = types.ModuleType(name)
        m.__dict__.update(attrs, _ca_stub=True)
        sys.modules[name] = m
        return m

    def _absent(*a, **k):
        raise RuntimeError("entmax is not installed (parity unpinned)")

    def have(name):
        try:
            __import__(name)
            return True
        except ImportError:
            return False

    if not have("entmax"):
        mod("entmax", entmax15=_absent, sparsemax=_absent)
    if not have("matplotlib"):
        mpl = mod("matplotlib")
        mpl.pyplot = mod("matplotlib.pyplot")
        mpl.animation = mod("matplotlib.animation")
    if not have("imwatermark"):
        class WatermarkEncoder:
            def set_watermark(self, *a, **k):
                pass

            def encode(self, x, *a, **k):
                return x
        mod("imwatermark", WatermarkEncoder=WatermarkEncoder)
    if not have("fire"):
        mod("fire", Fire=lambda *a, **k: None)


class _EinopsProxy:
    """einops with `rearrange(..., h=64, w=64)` of the '(h w)' unflatten replaced by isqrt(P)."""

    def __init__(self, real):
        self._real = real

    def __getattr__(self, name):
        return getattr(self._real, name)

    def rearrange(self, x, pattern, **axes):
        if "(h w)" in pattern and "h" in axes and "w" in axes:
            side = math.isqrt(x.shape[-1])
            if side * side == x.shape[-1]:
                axes["h"] = axes["w"] = side
        return self._real.rearrange(x, pattern, **axes)


def load():
    """-> namespace with the reference's hot-path classes and functions."""
    global _NS
    if _NS is not None:
        return _NS
    r = root()
    if r is None:
        raise RuntimeError("reference modules unavailable: neither oracle/_ref nor /root/reference exists "
                           "(run `python -m oracle.make_ref` in the build container)")
    install_stubs()
    if r not in sys.path:
        sys.path.insert(0, r)
    import concept_attention  # noqa: F401  (the reference's eager imports)
    from concept_attention import concept_attention_pipeline as pipeline
    from concept_attention import image_generator, modified_double_stream_block, modified_flux_dit
    from concept_attention import modified_single_stream_block, segmentation, utils
    from concept_attention.flux.src.flux import sampling
    from concept_attention.flux.src.flux.modules import autoencoder
    import einops
    pipeline.einops = _EinopsProxy(einops)          # patch (1)
    pipeline.print = lambda *a, **k: None           # patch (2)
    sampling.tqdm = lambda it, *a, **k: it          # patch (3)
    image_generator.print = lambda *a, **k: None
    ns = types.SimpleNamespace(
        root=r, kind="oracle/_ref" if r.endswith("_ref") else "/root/reference",
        pipeline=pipeline, sampling=sampling, segmentation=segmentation, utils=utils,
        image_generator=image_generator, autoencoder=autoencoder,
        ModifiedFluxDiT=modified_flux_dit.ModifiedFluxDiT, FluxParams=modified_flux_dit.FluxParams,
        ModifiedDoubleStreamBlock=modified_double_stream_block.ModifiedDoubleStreamBlock,
        ModifiedSingleStreamBlock=modified_single_stream_block.ModifiedSingleStreamBlock,
        compute_heatmaps_from_vectors=pipeline.compute_heatmaps_from_vectors,
        denoise=sampling.denoise, get_schedule=sampling.get_schedule, get_noise=sampling.get_noise,
        prepare=sampling.prepare, unpack=sampling.unpack, embed_concepts=utils.embed_concepts)
    _NS = ns
    return ns


def build_dit(params: dict, state_dict: dict, device, dtype):
    """Reference `ModifiedFluxDiT(FluxParams(**params))` holding `state_dict` (reference keys)
    cast to `dtype` on `device`; tensors already of that dtype/device are shared, not copied."""
    import torch
    ns = load()
    with torch.device("meta"):
        model = ns.ModifiedFluxDiT(ns.FluxParams(**params))
    sd = {k: v.to(device=device, dtype=dtype) for k, v in state_dict.items()}
    missing, unexpected = model.load_state_dict(sd, strict=True, assign=True)
    assert not missing and not unexpected, (missing, unexpected)
    for p in model.parameters():
        p.requires_grad_(False)
    return model.eval()


def heatmaps(stacked: dict, layer_indices, timesteps, **kw):
    """Both heatmap sets the reference pipeline derives from a denoise() dict
    (concept_attention_pipeline.py:150-169): (concept_heatmaps, cross_attention_maps)."""
    ns = load()
    out = ns.compute_heatmaps_from_vectors(
        stacked["output_space_image_vectors"], stacked["output_space_concept_vectors"],
        layer_indices=list(layer_indices), timesteps=list(timesteps), **kw)
    cross = ns.compute_heatmaps_from_vectors(
        stacked["cross_attention_image_vectors"], stacked["cross_attention_concept_vectors"],
        layer_indices=list(layer_indices), timesteps=list(timesteps), **kw)
    return out, cross
