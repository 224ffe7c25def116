"""Oracle heatmap maths and heatmap->mask steps (fp32, CPU).
TEST INFRASTRUCTURE — see oracle/__init__.py.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

__all__ = ["linear_normalization", "sparsemax", "entmax15", "heatmaps_from_vectors",
           "gaussian_blur3", "binary_masks", "argmax_mask", "multiclass_logits",
           "map_background_to_zero"]


def linear_normalization(x, dim):
    """min-shift then divide by the sum, zero sums -> 1 (utils.py:35-44)."""
    x = x - torch.min(x, dim=dim, keepdim=True)[0]
    s = torch.sum(x, dim=dim, keepdim=True)
    s = torch.where(s == 0, torch.ones_like(s), s)
    return x / s


def sparsemax(z, dim):
    """Euclidean projection onto the simplex (Martins & Astudillo 2016, Alg. 1).
    Parity UNPINNED: stands in for `entmax.sparsemax`
    (concept_attention_pipeline.py:5,68-69), a third-party package the reference
    neither vendors nor pins."""
    z = z.transpose(dim, -1)
    zs, _ = torch.sort(z, dim=-1, descending=True)
    k = torch.arange(1, z.shape[-1] + 1, dtype=z.dtype)
    cs = zs.cumsum(-1)
    support = (1 + k * zs) > cs
    ksup = support.sum(-1, keepdim=True)
    tau = (cs.gather(-1, ksup - 1) - 1) / ksup.to(z.dtype)
    return torch.clamp(z - tau, min=0).transpose(dim, -1)


def entmax15(z, dim):
    """1.5-entmax, exact sort-based threshold (Peters, Niculae & Martins 2019,
    Alg. 2).  Parity UNPINNED, see `sparsemax`."""
    z = z.transpose(dim, -1)
    z = (z - z.max(-1, keepdim=True)[0]) / 2
    zs, _ = torch.sort(z, dim=-1, descending=True)
    k = torch.arange(1, z.shape[-1] + 1, dtype=z.dtype)
    mean = zs.cumsum(-1) / k
    mean_sq = (zs * zs).cumsum(-1) / k
    ss = k * (mean_sq - mean * mean)
    delta = torch.clamp((1 - ss) / k, min=0)
    tau = mean - torch.sqrt(delta)
    ksup = (tau <= zs).sum(-1, keepdim=True)
    tau_star = tau.gather(-1, ksup - 1)
    return (torch.clamp(z - tau_star, min=0) ** 2).transpose(dim, -1)


def heatmaps_from_vectors(image_vectors, concept_vectors, layer_indices, timesteps=(0, 1, 2, 3),
                          softmax=True, normalize_concepts=False, attention_norm="sparsemax"):
    """compute_heatmaps_from_vectors (concept_attention_pipeline.py:29-91) with the
    hard-coded 64x64 (:85-90) replaced by isqrt(P) and the debug print (:72-74)
    dropped.  image [Nt,L,B,P,D] or [Nt,L,B,H,P,d]; returns [B,C,h,w]."""
    if image_vectors.ndim == 6:                         # collapse heads, :43-51
        def collapse(x):
            t, l, b, h, n, d = x.shape
            return x.permute(0, 1, 2, 4, 3, 5).reshape(t, l, b, n, h * d)
        image_vectors, concept_vectors = collapse(image_vectors), collapse(concept_vectors)
    if normalize_concepts:
        concept_vectors = linear_normalization(concept_vectors, dim=-2)
    maps = torch.einsum("tlbpd,tlbcd->tlbcp", image_vectors, concept_vectors)
    if softmax or attention_norm == "softmax":          # :64 — default path
        maps = torch.softmax(maps, dim=-2)
    elif attention_norm == "entmax15":
        maps = entmax15(maps, dim=-2)
    elif attention_norm == "sparsemax":
        maps = sparsemax(maps, dim=-2)
    else:
        raise ValueError(f"Unknown attention_norm={attention_norm}")
    maps = maps[list(timesteps)][:, list(layer_indices)]
    maps = maps.mean(dim=(0, 1))
    side = math.isqrt(maps.shape[-1])
    if side * side != maps.shape[-1]:                   # non-square token count: keep [B,C,P]
        return maps
    return maps.reshape(*maps.shape[:-1], side, side)


def gaussian_blur3(x, sigma=1.0):
    """torchvision.transforms.functional.gaussian_blur(kernel_size=3, sigma=1)
    as called at segmentation.py:57-58: separable kernel exp(-x^2/2s^2) on
    {-1,0,1}, normalised, reflect padding.  x [..., h, w]."""
    k = torch.exp(-torch.tensor([-1.0, 0.0, 1.0]) ** 2 / (2 * sigma * sigma))
    k = (k / k.sum()).to(x.dtype)
    k2 = torch.outer(k, k)[None, None]
    shp = x.shape
    y = F.pad(x.reshape(-1, 1, *shp[-2:]), (1, 1, 1, 1), mode="reflect")
    return F.conv2d(y, k2).reshape(shp)


def binary_masks(coefficients, apply_blur=False):
    """coeff_c > mean_hw(coeff_c) per concept (segmentation.py:56-61)."""
    if apply_blur:
        coefficients = gaussian_blur3(coefficients)
    return coefficients > coefficients.mean(dim=(-2, -1), keepdim=True)


def multiclass_logits(image_vectors, concept_vectors):
    """'patches dim, concepts dim -> concepts h w' without softmax
    (experiments/pascal_voc_segmentation/multi_class_segmentation.py:59-69)."""
    maps = torch.einsum("pd,cd->cp", image_vectors, concept_vectors)
    side = math.isqrt(maps.shape[-1])
    return maps.reshape(-1, side, side)


def argmax_mask(heatmaps, apply_blur=False):
    """argmax over the concept axis, first index on ties
    (multi_class_segmentation.py:71-77).  heatmaps [C,h,w] -> int64 [h,w]."""
    if apply_blur:
        heatmaps = gaussian_blur3(heatmaps)
    return heatmaps.argmax(dim=0)


def map_background_to_zero(pred, num_background):
    """predictions[predictions < len(background)] = 0
    (run_multi_class_seg_experiment.py:25-31)."""
    pred = pred.clone()
    pred[pred < num_background] = 0
    return pred
