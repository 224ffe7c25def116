"""Model specs and random-init weights with the reference's state-dict keys.

`FluxParams` mirrors concept_attention/modified_flux_dit.py:13-26 and `CONFIGS`
the table at concept_attention/flux/src/flux/util.py:28-93, so a real
`flux1-schnell.safetensors` state dict drops into `FluxWeights` unchanged.  No
checkpoint can be fetched offline, so `init_state_dict` draws the same
distribution `nn.Linear` uses by default (U(-1/sqrt(in), 1/sqrt(in)) for weight
and bias; RMSNorm scales = 1) from a seeded generator.
"""
from __future__ import annotations

import math
from dataclasses import asdict, dataclass, field

import torch

__all__ = ["FluxParams", "CONFIGS", "T5_MAX_LENGTH", "linear_shapes", "init_state_dict",
           "expected_keys", "load_checkpoint"]


@dataclass
class FluxParams:
    in_channels: int = 64
    vec_in_dim: int = 768
    context_in_dim: int = 4096
    hidden_size: int = 3072
    mlp_ratio: float = 4.0
    num_heads: int = 24
    depth: int = 19
    depth_single_blocks: int = 38
    axes_dim: list = field(default_factory=lambda: [16, 56, 56])
    theta: int = 10_000
    qkv_bias: bool = True
    guidance_embed: bool = False

    def asdict(self):
        return asdict(self)


CONFIGS = {
    "flux-schnell": FluxParams(guidance_embed=False),
    "flux-dev": FluxParams(guidance_embed=True),
    # 2 heads x 128: smallest shape the d=128 attention kernel accepts; used by
    # the golden fixtures and the fast tests.
    "tiny": FluxParams(in_channels=64, vec_in_dim=32, context_in_dim=64, hidden_size=256,
                       num_heads=2, depth=3, depth_single_blocks=2, guidance_embed=False),
    # full Flux width (D=3072, 24 heads) but 2 double + 1 single block: BASELINE config 1 shapes at a
    # depth the fp32 CPU oracle finishes in seconds (full-dimension parity test)
    "schnell-2l": FluxParams(depth=2, depth_single_blocks=1, guidance_embed=False),
    "tiny-dev": FluxParams(in_channels=64, vec_in_dim=32, context_in_dim=64, hidden_size=256,
                           num_heads=2, depth=2, depth_single_blocks=1, guidance_embed=True),
}

# T5 sequence length per model (concept_attention/image_generator.py:57)
T5_MAX_LENGTH = {"flux-schnell": 256, "flux-dev": 512, "tiny": 128, "tiny-dev": 128, "schnell-2l": 256}


def linear_shapes(p: FluxParams) -> dict:
    """{reference linear prefix: (out_features, in_features)} for every nn.Linear
    of ModifiedFluxDiT (modified_flux_dit.py:50-73 and the block ctors)."""
    D, M = p.hidden_size, int(p.hidden_size * p.mlp_ratio)
    s = {
        "img_in": (D, p.in_channels),
        "time_in.in_layer": (D, 256), "time_in.out_layer": (D, D),
        "vector_in.in_layer": (D, p.vec_in_dim), "vector_in.out_layer": (D, D),
        "txt_in": (D, p.context_in_dim),
        "final_layer.linear": (p.in_channels, D),
        "final_layer.adaLN_modulation.1": (2 * D, D),
    }
    if p.guidance_embed:
        s["guidance_in.in_layer"] = (D, 256)
        s["guidance_in.out_layer"] = (D, D)
    for i in range(p.depth):
        b = f"double_blocks.{i}"
        for st in ("img", "txt"):
            s[f"{b}.{st}_mod.lin"] = (6 * D, D)
            s[f"{b}.{st}_attn.qkv"] = (3 * D, D)
            s[f"{b}.{st}_attn.proj"] = (D, D)
            s[f"{b}.{st}_mlp.0"] = (M, D)
            s[f"{b}.{st}_mlp.2"] = (D, M)
    for i in range(p.depth_single_blocks):
        b = f"single_blocks.{i}"
        s[f"{b}.linear1"] = (3 * D + M, D)
        s[f"{b}.linear2"] = (D, D + M)
        s[f"{b}.modulation.lin"] = (3 * D, D)
    return s


def init_state_dict(p: FluxParams, seed: int = 0, device="cpu", dtype=torch.float32,
                    norm_scale_jitter: float = 0.0) -> dict:
    """Seeded random-init state dict (reference key names).

    Values are drawn on `device` in fp32 then cast; with device='cpu' the result
    is bit-reproducible across machines for a given torch version, which is what
    the golden fixtures rely on.  `norm_scale_jitter` > 0 perturbs the RMSNorm
    scales away from 1 so tests exercise that multiply.
    """
    g = torch.Generator(device=device).manual_seed(seed)
    sd = {}

    def uniform(shape, bound):
        t = torch.rand(shape, generator=g, device=device, dtype=torch.float32)
        return ((t * 2 - 1) * bound).to(dtype)

    for name, (out_f, in_f) in linear_shapes(p).items():
        bound = 1.0 / math.sqrt(in_f)
        sd[name + ".weight"] = uniform((out_f, in_f), bound)
        if p.qkv_bias or not name.endswith("_attn.qkv"):
            sd[name + ".bias"] = uniform((out_f,), bound)
    d = p.hidden_size // p.num_heads
    norm_keys = [f"double_blocks.{i}.{st}_attn.norm.{qk}_norm.scale"
                 for i in range(p.depth) for st in ("img", "txt") for qk in ("query", "key")]
    norm_keys += [f"single_blocks.{i}.norm.{qk}_norm.scale"
                  for i in range(p.depth_single_blocks) for qk in ("query", "key")]
    for k in norm_keys:
        s = torch.ones(d, device=device, dtype=torch.float32)
        if norm_scale_jitter:
            s = s + norm_scale_jitter * (torch.rand(d, generator=g, device=device) * 2 - 1)
        sd[k] = s.to(dtype)
    return sd


def expected_keys(p: FluxParams) -> dict:
    """{state-dict key: shape} of ModifiedFluxDiT for `p` (what a flux1-*.safetensors holds)."""
    shapes = {}
    for name, (out_f, in_f) in linear_shapes(p).items():
        shapes[name + ".weight"] = (out_f, in_f)
        if p.qkv_bias or not name.endswith("_attn.qkv"):
            shapes[name + ".bias"] = (out_f,)
    d = p.hidden_size // p.num_heads
    for i in range(p.depth):
        for st in ("img", "txt"):
            for qk in ("query", "key"):
                shapes[f"double_blocks.{i}.{st}_attn.norm.{qk}_norm.scale"] = (d,)
    for i in range(p.depth_single_blocks):
        for qk in ("query", "key"):
            shapes[f"single_blocks.{i}.norm.{qk}_norm.scale"] = (d,)
    return shapes


def load_checkpoint(path: str, p: FluxParams, device="cpu", dtype=torch.bfloat16, strict=True) -> dict:
    """Read a Flux checkpoint with the reference's key names — the file
    `load_flow_model` feeds to `model.load_state_dict` (concept_attention/image_generator.py:19-48): `.safetensors` (tensors are read one by one straight to
    `device`) or a torch `.pt/.pth` state dict.  Keys and shapes are checked against
    `expected_keys(p)`; `strict=False` tolerates extra keys (e.g. the reference's
    non-persistent buffers) but never missing or mis-shaped ones."""
    want = expected_keys(p)
    sd = {}
    if str(path).endswith(".safetensors"):
        from safetensors import safe_open
        with safe_open(str(path), framework="pt", device=str(device)) as f:
            names = list(f.keys())
            for k in names:
                if k in want:
                    sd[k] = f.get_tensor(k).to(dtype)
    else:
        raw = torch.load(str(path), map_location=device, weights_only=True)
        names = list(raw.keys())
        sd = {k: v.to(dtype) for k, v in raw.items() if k in want}
    missing = sorted(set(want) - set(sd))
    extra = sorted(set(names) - set(want))
    bad = sorted(k for k, v in sd.items() if tuple(v.shape) != tuple(want[k]))
    if missing or bad or (strict and extra):
        raise ValueError(f"checkpoint does not match {type(p).__name__}: missing={missing[:5]} "
                         f"({len(missing)}), wrong shape={bad[:5]} ({len(bad)}), unexpected={extra[:5]} ({len(extra)})")
    return sd
