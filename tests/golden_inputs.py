"""Seed-derived inputs shared by tests/golden/make_golden.py (which feeds them to
the reference) and the tests (which feed them to the oracle / the CUDA path)."""
import math

import torch

WEIGHT_SEED = 11
INPUT_SEED = 23
TINY_SIDE = 192        # -> P = 144 image tokens: not a multiple of the 128-row tiles
TINY_T = 72
TINY_C = 3
HM_LAYERS = [1, 2]
HM_TIMESTEPS = [0, 1]
ABLATIONS = {
    "cross_only": {"concept_cross_attention": True, "concept_self_attention": False},
    "self_only": {"concept_cross_attention": False, "concept_self_attention": True},
    "neither": {"concept_cross_attention": False, "concept_self_attention": False},
    "both": {"concept_cross_attention": True, "concept_self_attention": True},
}


def dit_inputs(p, side, T, C, seed, batch=1, concept_scale=1.0):
    """Synthetic stand-ins for prepare()/embed_concepts() outputs
    (flux/src/flux/sampling.py:31-65, utils.py:6-33): patchified noise latents,
    T5-like txt, CLIP-like vec, first-token concept embeddings, zero ids and a
    zeroed concept CLIP vector."""
    g = torch.Generator().manual_seed(seed)
    h2 = w2 = math.ceil(side / 16)
    lat = torch.randn(batch, 16, 2 * h2, 2 * w2, generator=g)
    img = lat.reshape(batch, 16, h2, 2, w2, 2).permute(0, 2, 4, 1, 3, 5).reshape(batch, h2 * w2, 64)
    ids = torch.zeros(h2, w2, 3)
    ids[..., 1] = torch.arange(h2)[:, None].float()
    ids[..., 2] = torch.arange(w2)[None, :].float()
    return {
        "img": img,
        "img_ids": ids.reshape(1, -1, 3).repeat(batch, 1, 1),
        "txt": torch.randn(batch, T, p.context_in_dim, generator=g),
        "txt_ids": torch.zeros(batch, T, 3),
        "vec": torch.randn(batch, p.vec_in_dim, generator=g),
        "concepts": torch.randn(batch, C, p.context_in_dim, generator=g) * concept_scale,
        "concept_ids": torch.zeros(batch, C, 3),
        "concept_vec": torch.zeros(batch, p.vec_in_dim),
    }


def heatmap_inputs():
    """Vector stacks for compute_heatmaps_from_vectors at P=4096 (the only size
    the unmodified reference function accepts).  'out' is the 5-D output-space
    form, 'cross' the 6-D per-head form."""
    g = torch.Generator().manual_seed(77)
    Nt, L, B, P, C = 2, 3, 1, 4096, 5
    out = (torch.randn(Nt, L, B, P, 48, generator=g), torch.randn(Nt, L, B, C, 48, generator=g) * 0.6)
    cross = (torch.randn(Nt, L, B, 2, P, 24, generator=g), torch.randn(Nt, L, B, 2, C, 24, generator=g) * 0.6)
    return {"out": out, "cross": cross}


def mask_inputs():
    """[5,16,16] coefficients quantised to 1/4 so the argmax has many exact ties."""
    g = torch.Generator().manual_seed(99)
    return torch.round(torch.randn(5, 16, 16, generator=g) * 2) / 4


def metric_inputs():
    """Seeded label maps / scores for the segmentation metrics (utils.py:48-108).

    binary: the (1 - mask, mask) float stacks of experiments/imagenet_segmentation/run_experiment.py:206-214
    voc:    integer labels 0..20 with the ignore label -1 (nclass = 21)
    ap:     scores quantised to 1/16 (many ties) + labels with some ignored pixels
    """
    g = torch.Generator().manual_seed(1234)
    m = (torch.rand(224, 224, generator=g) > 0.6).float()
    t = (torch.rand(224, 224, generator=g) > 0.5).float()
    binary = (torch.stack((1 - m, m)), torch.stack((1 - t, t)), 2)
    vp = torch.randint(0, 21, (3, 64, 64), generator=g).float()
    vt = torch.randint(-1, 21, (3, 64, 64), generator=g).float()
    vp = torch.where(torch.rand(3, 64, 64, generator=g) > 0.5, vt.clamp(min=0), vp)   # ~half correct
    voc = (vp, vt, 21)
    c = torch.round(torch.rand(2, 64, 64, generator=g) * 16) / 16
    lab = (torch.rand(2, 64, 64, generator=g) > 0.7).float()
    lab[0, :4] = -1.0                                                                # ignored rows
    ap = (torch.stack((1 - c, c), dim=1), lab)                                       # [2,2,64,64], [2,64,64]
    return {"binary": binary, "voc": voc, "ap": ap}
