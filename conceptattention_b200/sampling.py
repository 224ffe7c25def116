"""Device-side sampling helpers and synthetic text encoders.

Mirrors concept_attention/flux/src/flux/sampling.py (get_noise :12-29, prepare
:31-65, get_schedule :78-94, unpack :154-162), concept_attention/utils.py
embed_concepts (:6-33) and segmentation.add_noise_to_image (:85-113).  T5-XXL /
CLIP-L weights cannot be fetched offline, so `SyntheticEmbedder` stands in for
`HFEmbedder` (flux/modules/conditioner.py:6-38): a deterministic, prompt-seeded
embedding of the same shape.  A real HFEmbedder instance can be passed instead.
"""
from __future__ import annotations

import math
import zlib

import torch


class SyntheticEmbedder:
    """Callable like HFEmbedder: list[str] | str -> [n, max_length, dim] (T5-like)
    or [n, dim] (CLIP-like, pooled).  Values are randn seeded by crc32(prompt)."""

    def __init__(self, dim: int, max_length: int | None, device="cpu", dtype=torch.bfloat16, pin=False):
        """`pin=True` (host embedder feeding a GPU): per-prompt embeddings and the stacked result
        live in pinned host memory, so the `.to(cuda, non_blocking=True)` that follows is a real
        asynchronous H2D copy instead of a staged pageable one."""
        self.dim, self.max_length, self.device, self.dtype = dim, max_length, device, dtype
        self.pin = pin and torch.cuda.is_available()
        self._cache = {}

    def to(self, device):
        self.device = device
        return self

    def cpu(self):
        return self.to("cpu")

    def __call__(self, text):
        if isinstance(text, str):
            text = [text]
        outs = []
        for t in text:
            if t not in self._cache:
                g = torch.Generator().manual_seed(zlib.crc32(t.encode("utf-8")))
                shape = (self.dim,) if self.max_length is None else (self.max_length, self.dim)
                e = torch.randn(shape, generator=g).to(self.dtype)
                self._cache[t] = e.pin_memory() if self.pin else e
            outs.append(self._cache[t])
        if len(outs) == 1:
            return outs[0][None].to(self.device)            # a view: stays pinned
        if self.pin:
            out = torch.empty((len(outs),) + tuple(outs[0].shape), dtype=self.dtype, pin_memory=True)
            return torch.stack(outs, out=out).to(self.device)
        return torch.stack(outs).to(self.device)


def get_noise(num_samples, height, width, device, dtype, seed):
    return torch.randn(num_samples, 16, 2 * math.ceil(height / 16), 2 * math.ceil(width / 16),
                       device=device, dtype=dtype,
                       generator=torch.Generator(device=device).manual_seed(seed))


def patchify(lat):
    """'b c (h 2) (w 2) -> b (h w) (c 2 2)'"""
    b, c, H, W = lat.shape
    return (lat.reshape(b, c, H // 2, 2, W // 2, 2).permute(0, 2, 4, 1, 3, 5)
            .reshape(b, (H // 2) * (W // 2), c * 4))


def unpack(x, height, width):
    h, w = math.ceil(height / 16), math.ceil(width / 16)
    return (x.reshape(x.shape[0], h, w, -1, 2, 2).permute(0, 3, 1, 4, 2, 5)
            .reshape(x.shape[0], -1, h * 2, w * 2))


def image_ids(batch, h2, w2, device):
    ids = torch.zeros(h2, w2, 3, device=device)
    ids[..., 1] = torch.arange(h2, device=device)[:, None].float()
    ids[..., 2] = torch.arange(w2, device=device)[None, :].float()
    return ids.reshape(1, h2 * w2, 3).repeat(batch, 1, 1)


def prepare(t5, clip, img, prompt):
    """Same dict as the reference prepare(): img, img_ids, txt, txt_ids, vec."""
    bs, c, h, w = img.shape
    if bs == 1 and not isinstance(prompt, str):
        bs = len(prompt)
    x = patchify(img)
    if x.shape[0] == 1 and bs > 1:
        x = x.repeat(bs, 1, 1)
    if isinstance(prompt, str):
        prompt = [prompt]
    txt = t5(prompt)
    if txt.shape[0] == 1 and bs > 1:
        txt = txt.repeat(bs, 1, 1)
    vec = clip(prompt)
    if vec.shape[0] == 1 and bs > 1:
        vec = vec.repeat(bs, 1)
    dev = img.device
    return {"img": x, "img_ids": image_ids(bs, h // 2, w // 2, dev), "txt": txt.to(dev, non_blocking=True),
            "txt_ids": torch.zeros(bs, txt.shape[1], 3, device=dev), "vec": vec.to(dev, non_blocking=True)}


def embed_concepts(clip, t5, concepts, batch_size=1):
    """First T5 token per concept, zero ids, zeroed CLIP vector (utils.py:6-33)."""
    toks = [t5(c)[0, 0, :] for c in concepts]
    if toks[0].device.type == "cpu" and getattr(t5, "pin", False):
        emb = torch.stack(toks, out=torch.empty((len(toks),) + tuple(toks[0].shape), dtype=toks[0].dtype,
                                                pin_memory=True)).unsqueeze(0)
    else:
        emb = torch.stack(toks).unsqueeze(0)
    ids = torch.zeros(batch_size, emb.shape[1], 3)
    vec = torch.zeros_like(clip(" ".join(concepts)))
    return emb, ids, vec


def get_schedule(num_steps, image_seq_len, base_shift=0.5, max_shift=1.15, shift=True):
    ts = torch.linspace(1, 0, num_steps + 1)
    if shift:
        m = (max_shift - base_shift) / (4096 - 256)
        mu = m * image_seq_len + (base_shift - m * 256)
        ts = math.exp(mu) / (math.exp(mu) + (1 / ts - 1) ** 1.0)
    return ts.tolist()


def add_noise_to_image(encoded_image, num_steps=50, noise_timestep=49, seed=63, width=1024,
                       height=1024, device="cuda", is_schnell=True):
    x = get_noise(1, height, width, device=device, dtype=torch.bfloat16, seed=seed)
    ts = get_schedule(num_steps, x.shape[-1] * x.shape[-2] // 4, shift=(not is_schnell))
    t = ts[noise_timestep]
    return t * x + (1.0 - t) * encoded_image.to(x.dtype), ts[noise_timestep:]


# plasma colormap.  With matplotlib importable its own 256-entry table is used (exact reference
# behaviour).  Without it (this image has no matplotlib and no network to fetch the table from) a
# 256-entry LUT is built once from 10 anchor colours of the matplotlib table by linear
# interpolation and LOOKED UP the way matplotlib does it (Colormap.__call__): index =
# int(x * 256) clipped to 255, no interpolation between entries, NaN -> the "bad" colour (0, 0, 0).
_PLASMA = [(13, 8, 135), (70, 3, 159), (114, 1, 168), (156, 23, 158), (189, 55, 134),
           (216, 87, 107), (237, 121, 83), (251, 159, 58), (253, 202, 38), (240, 249, 33)]
_PLASMA_LUT = None


def plasma_lut():
    """uint8 [256,3] stand-in for matplotlib's plasma table (anchors exact, in between linear)."""
    global _PLASMA_LUT
    if _PLASMA_LUT is None:
        import numpy as np
        anchors = np.asarray(_PLASMA, dtype=np.float64)
        pos = np.arange(256) / 255.0 * (len(anchors) - 1)
        lo = np.floor(pos).astype(int).clip(0, len(anchors) - 2)
        f = (pos - lo)[:, None]
        _PLASMA_LUT = np.rint(anchors[lo] * (1 - f) + anchors[lo + 1] * f).astype(np.uint8)
    return _PLASMA_LUT


def apply_colormap(x01, cmap="plasma"):
    """[h,w] in [0,1] -> uint8 [h,w,3] (concept_attention_pipeline.py:180-183)."""
    import numpy as np
    try:
        import matplotlib.pyplot as plt
        return (plt.get_cmap(cmap)(x01)[:, :, :3] * 255).astype(np.uint8)
    except (ImportError, AttributeError):       # absent, or a stub module someone put in sys.modules
        if cmap != "plasma":
            raise ValueError(f"cmap={cmap!r} needs matplotlib; only 'plasma' is built in")
        x = np.asarray(x01, dtype=np.float64)
        bad = np.isnan(x)
        idx = np.clip((np.nan_to_num(x) * 256).astype(np.int64), 0, 255)
        out = plasma_lut()[idx]
        out[bad] = 0
        return out
