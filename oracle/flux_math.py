"""Oracle primitives (fp32, CPU).  TEST INFRASTRUCTURE — see oracle/__init__.py.

Each function names the reference lines it restates (paths under
/root/reference/concept_attention/).
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

__all__ = [
    "rope_table", "apply_rope", "rms_norm_qk", "layer_norm_modulate", "modulation",
    "timestep_embedding", "mlp_embedder", "linear", "split_heads", "merge_heads",
    "sdpa", "sdpa_explicit", "get_noise", "patchify", "unpatchify", "image_ids",
    "get_schedule", "gelu_tanh",
]


def linear(sd: dict, prefix: str, x: torch.Tensor) -> torch.Tensor:
    """nn.Linear with weights looked up by reference state-dict key."""
    return F.linear(x, sd[prefix + ".weight"], sd.get(prefix + ".bias"))


def rope_table(ids: torch.Tensor, axes_dim, theta: int) -> torch.Tensor:
    """ids [B,N,A] -> rotation matrices [B,1,N,sum(axes)/2,2,2] fp32.

    Restates flux/src/flux/math.py:15-22 (`rope`: fp64 angles
    pos * theta^(-2i/dim), matrix [[cos,-sin],[sin,cos]], cast to fp32) and
    flux/src/flux/modules/layers.py:11-25 (`EmbedND`: concat per axis, add a
    head axis).
    """
    mats = []
    for a, dim in enumerate(axes_dim):
        assert dim % 2 == 0
        expo = torch.arange(0, dim, 2, dtype=torch.float64, device=ids.device) / dim
        omega = 1.0 / (float(theta) ** expo)
        ang = ids[..., a].to(torch.float64)[..., None] * omega      # [B,N,dim/2]
        c, s = torch.cos(ang), torch.sin(ang)
        m = torch.stack([c, -s, s, c], dim=-1).reshape(*ang.shape, 2, 2)
        mats.append(m.float())
    return torch.cat(mats, dim=-3).unsqueeze(1)


def apply_rope(q: torch.Tensor, k: torch.Tensor, pe: torch.Tensor):
    """Pairwise 2x2 rotation in fp32, cast back (flux/src/flux/math.py:25-30)."""
    def rot(x):
        xf = x.float().reshape(*x.shape[:-1], -1, 1, 2)
        out = pe[..., 0] * xf[..., 0] + pe[..., 1] * xf[..., 1]
        return out.reshape(*x.shape).type_as(x)
    return rot(q), rot(k)


def rms_norm_qk(q, k, v, q_scale, k_scale):
    """QKNorm (layers.py:63-84): fp32 x*rsqrt(mean(x^2)+1e-6), cast to the input
    dtype, times the learned per-channel scale, then `.to(v)`."""
    def rms(x, scale):
        xf = x.float()
        r = torch.rsqrt(torch.mean(xf * xf, dim=-1, keepdim=True) + 1e-6)
        return (xf * r).to(x.dtype) * scale
    return rms(q, q_scale).to(v), rms(k, k_scale).to(v)


def layer_norm_modulate(x, shift, scale):
    """(1+scale)*LayerNorm(x, eps=1e-6, no affine)+shift
    (modified_double_stream_block.py:88-89,94-95,100-101)."""
    xn = F.layer_norm(x, (x.shape[-1],), eps=1e-6)
    return (1 + scale) * xn + shift


def modulation(sd, prefix, vec, n_chunks):
    """Modulation.forward (layers.py:113-126): Linear(SiLU(vec))[:,None,:]
    chunked into `n_chunks` (shift,scale,gate[,shift2,scale2,gate2])."""
    out = linear(sd, prefix + ".lin", F.silu(vec))[:, None, :]
    return out.chunk(n_chunks, dim=-1)


def timestep_embedding(t, dim, max_period=10000, time_factor=1000.0):
    """Sinusoidal embedding (layers.py:28-49), cos first then sin."""
    t = time_factor * t
    half = dim // 2
    freqs = torch.exp(-math.log(max_period) * torch.arange(half, dtype=torch.float32, device=t.device) / half)
    args = t[:, None].float() * freqs[None]
    emb = torch.cat([torch.cos(args), torch.sin(args)], dim=-1)
    if dim % 2:
        emb = torch.cat([emb, torch.zeros_like(emb[:, :1])], dim=-1)
    if torch.is_floating_point(t):
        emb = emb.to(t)
    return emb


def mlp_embedder(sd, prefix, x):
    """MLPEmbedder (layers.py:52-60): out_layer(SiLU(in_layer(x)))."""
    return linear(sd, prefix + ".out_layer", F.silu(linear(sd, prefix + ".in_layer", x)))


def gelu_tanh(x):
    return F.gelu(x, approximate="tanh")


def split_heads(qkv, num_heads):
    """'B L (K H D) -> K B H L D', K=3 (modified_double_stream_block.py:91)."""
    B, L, _ = qkv.shape
    x = qkv.reshape(B, L, 3, num_heads, -1).permute(2, 0, 3, 1, 4)
    return x[0], x[1], x[2]


def merge_heads(x):
    """'B H L D -> B L (H D)' (modified_double_stream_block.py:170-176)."""
    B, H, L, D = x.shape
    return x.permute(0, 2, 1, 3).reshape(B, L, H * D)


def sdpa(q, k, v):
    """softmax(q k^T / sqrt(d)) v, no mask — the library call the reference
    makes (modified_double_stream_block.py:112-116,162-166)."""
    return F.scaled_dot_product_attention(q, k, v)


def sdpa_explicit(q, k, v):
    """Explicit-matrix attention used by the ablation branches
    (modified_double_stream_block.py:21-41)."""
    w = q @ k.transpose(-2, -1) * (1.0 / math.sqrt(q.shape[-1]))
    return torch.softmax(w, dim=-1) @ v


def get_noise(num_samples, height, width, dtype, seed):
    """flux/src/flux/sampling.py:12-29, CPU generator."""
    return torch.randn(
        num_samples, 16, 2 * math.ceil(height / 16), 2 * math.ceil(width / 16),
        dtype=dtype, generator=torch.Generator().manual_seed(seed))


def patchify(lat):
    """'b c (h ph) (w pw) -> b (h w) (c ph pw)', ph=pw=2 (sampling.py:36)."""
    b, c, H, W = lat.shape
    x = lat.reshape(b, c, H // 2, 2, W // 2, 2).permute(0, 2, 4, 1, 3, 5)
    return x.reshape(b, (H // 2) * (W // 2), c * 4)


def unpatchify(x, height, width):
    """Inverse of patchify (sampling.py:154-162)."""
    h, w = math.ceil(height / 16), math.ceil(width / 16)
    b = x.shape[0]
    y = x.reshape(b, h, w, -1, 2, 2).permute(0, 3, 1, 4, 2, 5)
    return y.reshape(b, -1, h * 2, w * 2)


def image_ids(batch, h2, w2):
    """img_ids[...,1]=row, [...,2]=col, [...,0]=0 (sampling.py:40-43)."""
    ids = torch.zeros(h2, w2, 3)
    ids[..., 1] = torch.arange(h2)[:, None].float()
    ids[..., 2] = torch.arange(w2)[None, :].float()
    return ids.reshape(1, h2 * w2, 3).repeat(batch, 1, 1)


def get_schedule(num_steps, image_seq_len, base_shift=0.5, max_shift=1.15, shift=True):
    """linspace(1,0,n+1), optional resolution-dependent time shift
    (sampling.py:67-94)."""
    ts = torch.linspace(1, 0, num_steps + 1)
    if shift:
        m = (max_shift - base_shift) / (4096 - 256)
        mu = m * image_seq_len + (base_shift - m * 256)
        ts = math.exp(mu) / (math.exp(mu) + (1 / ts - 1) ** 1.0)
    return ts.tolist()
