"""Oracle DiT blocks (fp32, CPU).  TEST INFRASTRUCTURE — see oracle/__init__.py.

Functional restatement of the reference modules; weights come from a state dict
with the reference's own keys (`double_blocks.{i}.img_attn.qkv.weight`, ...), so
the same dict can be loaded into the reference `ModifiedFluxDiT` for pinning.
"""
from __future__ import annotations

import torch

from .flux_math import (apply_rope, gelu_tanh, layer_norm_modulate, linear, merge_heads,
                        mlp_embedder, modulation, rms_norm_qk, rope_table, sdpa,
                        sdpa_explicit, split_heads, timestep_embedding)

__all__ = ["double_block", "single_block", "dit_forward", "denoise", "VECTOR_KEYS"]

VECTOR_KEYS = (
    "output_space_concept_vectors", "output_space_image_vectors",
    "cross_attention_concept_vectors", "cross_attention_image_vectors",
)


def _stream_qkv(sd, p, stream, x, shift, scale, H):
    """LN+modulate -> qkv Linear -> split heads -> QK RMSNorm for one stream.
    `stream` names the weight set ('img' or 'txt'); concepts pass 'txt'
    (modified_double_stream_block.py:88-104)."""
    xm = layer_norm_modulate(x, shift, scale)
    q, k, v = split_heads(linear(sd, f"{p}.{stream}_attn.qkv", xm), H)
    q, k = rms_norm_qk(q, k, v, sd[f"{p}.{stream}_attn.norm.query_norm.scale"],
                       sd[f"{p}.{stream}_attn.norm.key_norm.scale"])
    return q, k, v


def _stream_update(sd, p, stream, x, attn, mod):
    """x += gate1*proj(attn); x += gate2*mlp((1+scale2)*LN(x)+shift2)
    (modified_double_stream_block.py:194-202)."""
    sh1, sc1, g1, sh2, sc2, g2 = mod
    x = x + g1 * linear(sd, f"{p}.{stream}_attn.proj", attn)
    h = layer_norm_modulate(x, sh2, sc2)
    h = linear(sd, f"{p}.{stream}_mlp.2", gelu_tanh(linear(sd, f"{p}.{stream}_mlp.0", h)))
    return x + g2 * h


def double_block(sd, p, num_heads, img, txt, vec, pe, concepts, concept_vec, concept_pe,
                 joint_attention_kwargs=None):
    """ModifiedDoubleStreamBlock.forward (modified_double_stream_block.py:69-204).

    Returns (img, txt, concepts, dict) with the four vector sets the reference
    collects at :185-191.  Like the reference this executes the full
    (C+P)x(C+P) second attention and keeps rows [:C].
    """
    assert concept_vec is not None
    img_mod = modulation(sd, f"{p}.img_mod", vec, 6)
    txt_mod = modulation(sd, f"{p}.txt_mod", vec, 6)
    con_mod = modulation(sd, f"{p}.txt_mod", concept_vec, 6)      # txt weights, :86

    iq, ik, iv = _stream_qkv(sd, p, "img", img, img_mod[0], img_mod[1], num_heads)
    tq, tk, tv = _stream_qkv(sd, p, "txt", txt, txt_mod[0], txt_mod[1], num_heads)
    cq, ck, cv = _stream_qkv(sd, p, "txt", concepts, con_mod[0], con_mod[1], num_heads)

    T, C = txt.shape[1], concepts.shape[1]
    # text (+) image joint attention, :106-119
    q, k = apply_rope(torch.cat((tq, iq), 2), torch.cat((tk, ik), 2), pe)
    ti = sdpa(q, k, torch.cat((tv, iv), 2))
    txt_attn, img_attn = ti[:, :, :T], ti[:, :, T:]
    # concept (+) image attention, :121-168
    q2, k2 = apply_rope(torch.cat((cq, iq), 2), torch.cat((ck, ik), 2), concept_pe)
    v2 = torch.cat((cv, iv), 2)
    cross = self_ = True
    if joint_attention_kwargs is not None:
        cross = joint_attention_kwargs.get("concept_cross_attention", True)
        self_ = joint_attention_kwargs.get("concept_self_attention", True)
    if cross and self_:
        concept_attn = sdpa(q2, k2, v2)[:, :, :C]
    elif cross:
        concept_attn = sdpa_explicit(q2[:, :, :C], k2[:, :, C:], iv)
    elif self_:
        concept_attn = sdpa_explicit(q2[:, :, :C], k2[:, :, :C], cv)
    else:
        concept_attn = cv

    txt_attn, img_attn, concept_attn = map(merge_heads, (txt_attn, img_attn, concept_attn))
    vectors = {
        "output_space_concept_vectors": concept_attn,
        "output_space_image_vectors": img_attn,
        "cross_attention_concept_vectors": cq,          # post-RMSNorm, pre-RoPE, :189
        "cross_attention_image_vectors": iq,            # post-RMSNorm, pre-RoPE, :190
    }
    img = _stream_update(sd, p, "img", img, img_attn, img_mod)
    txt = _stream_update(sd, p, "txt", txt, txt_attn, txt_mod)
    concepts = _stream_update(sd, p, "txt", concepts, concept_attn, con_mod)
    return img, txt, concepts, vectors


def single_block(sd, p, num_heads, x, vec, pe):
    """ModifiedSingleStreamBlock.forward (modified_single_stream_block.py:43-56)."""
    shift, scale, gate = modulation(sd, f"{p}.modulation", vec, 3)
    D = x.shape[-1]
    h = linear(sd, f"{p}.linear1", layer_norm_modulate(x, shift, scale))
    qkv, mlp = h[..., :3 * D], h[..., 3 * D:]
    q, k, v = split_heads(qkv, num_heads)
    q, k = rms_norm_qk(q, k, v, sd[f"{p}.norm.query_norm.scale"], sd[f"{p}.norm.key_norm.scale"])
    q, k = apply_rope(q, k, pe)
    attn = merge_heads(sdpa(q, k, v))
    out = linear(sd, f"{p}.linear2", torch.cat((attn, gelu_tanh(mlp)), 2))
    return x + gate * out


def dit_forward(sd, params, img, img_ids, txt, txt_ids, concepts, concept_ids, concept_vec,
                timesteps, y, guidance=None, stop_after_multimodal_attentions=False,
                joint_attention_kwargs=None):
    """ModifiedFluxDiT.forward (modified_flux_dit.py:75-163).

    `params` is a dict with hidden_size, num_heads, depth, depth_single_blocks,
    axes_dim, theta, guidance_embed.  Returns (pred | None, {key: [L,B,...]}).
    """
    assert concept_vec is not None
    if img.ndim != 3 or txt.ndim != 3:
        raise ValueError("Input img and txt tensors must have 3 dimensions.")
    H = params["num_heads"]

    def cond_vec(pooled):
        v = mlp_embedder(sd, "time_in", timestep_embedding(timesteps, 256))
        if params["guidance_embed"]:
            if guidance is None:
                raise ValueError("Didn't get guidance strength for guidance distilled model.")
            v = v + mlp_embedder(sd, "guidance_in", timestep_embedding(guidance, 256))
        return v + mlp_embedder(sd, "vector_in", pooled)

    img = linear(sd, "img_in", img)
    vec = cond_vec(y)
    txt = linear(sd, "txt_in", txt)
    pe = rope_table(torch.cat((txt_ids, img_ids), 1), params["axes_dim"], params["theta"])
    pe_c = rope_table(torch.cat((concept_ids, img_ids), 1), params["axes_dim"], params["theta"])
    cvec = cond_vec(concept_vec)                                   # :113-119
    concepts = linear(sd, "txt_in", concepts)                      # :120

    collected = {k: [] for k in VECTOR_KEYS}
    for i in range(params["depth"]):
        img, txt, concepts, vecs = double_block(
            sd, f"double_blocks.{i}", H, img, txt, vec, pe, concepts, cvec, pe_c,
            joint_attention_kwargs)
        for k in VECTOR_KEYS:
            collected[k].append(vecs[k])
    collected = {k: torch.stack(v, 0) for k, v in collected.items()}
    if stop_after_multimodal_attentions:
        return None, collected

    x = torch.cat((txt, img), 1)
    for i in range(params["depth_single_blocks"]):
        x = single_block(sd, f"single_blocks.{i}", H, x, vec, pe)
    x = x[:, txt.shape[1]:]
    shift, scale = linear(sd, "final_layer.adaLN_modulation.1",
                          torch.nn.functional.silu(vec)).chunk(2, dim=1)
    x = layer_norm_modulate(x, shift[:, None, :], scale[:, None, :])
    return linear(sd, "final_layer.linear", x), collected


def denoise(sd, params, img, img_ids, txt, txt_ids, vec, timesteps, guidance=4.0,
            concepts=None, concept_ids=None, concept_vec=None, joint_attention_kwargs=None):
    """Euler sampler that stacks every step's vectors (sampling.py:96-152).
    Returns (img, {key: [Nt,L,B,...]})."""
    collected = {k: [] for k in VECTOR_KEYS}
    g = torch.full((img.shape[0],), guidance, dtype=img.dtype, device=img.device)
    for t_curr, t_prev in zip(timesteps[:-1], timesteps[1:]):
        t_vec = torch.full((img.shape[0],), t_curr, dtype=img.dtype, device=img.device)
        pred, vecs = dit_forward(sd, params, img, img_ids, txt, txt_ids, concepts, concept_ids,
                                 concept_vec, t_vec, vec, g,
                                 joint_attention_kwargs=joint_attention_kwargs)
        img = img + (t_prev - t_curr) * pred
        for k in VECTOR_KEYS:
            collected[k].append(vecs[k])
    return img, {k: torch.stack(v, 0) for k, v in collected.items()}
