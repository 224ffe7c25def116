"""Host post-processing of the heatmaps (concept_attention_pipeline.py:171-196): ONE global
min-max over all concepts of an image, then a 256-entry colormap LOOKUP (matplotlib's
Colormap.__call__: index = int(x * 256) clipped to 255, no interpolation), RGB uint8, PIL images.
matplotlib is absent from this image, so the table itself is the built-in stand-in (anchors of
the matplotlib plasma table, linear in between); what is pinned here is the reference's semantics."""
import numpy as np

from conceptattention_b200 import sampling
from conceptattention_b200.pipeline import _colour


def test_plasma_lut_has_256_entries_and_the_matplotlib_end_points():
    lut = sampling.plasma_lut()
    assert lut.shape == (256, 3) and lut.dtype == np.uint8
    assert tuple(lut[0]) == (13, 8, 135) and tuple(lut[255]) == (240, 249, 33)   # plasma(0.0), plasma(1.0)
    assert tuple(lut[85]) == (156, 23, 158)                                      # an anchor on the grid: 85 = 3 * 255 / 9
    lum = lut.astype(np.float64) @ np.array([0.2126, 0.7152, 0.0722])
    assert (np.diff(lum) > -0.5).all()                                           # perceptually increasing map


def test_lookup_quantises_like_matplotlib():
    x = np.array([[0.0, 0.5 / 256, 0.999 / 256, 1.0 / 256, 0.5, 255.0 / 256, 1.0, 1.5, -0.2, np.nan]])
    rgb = sampling.apply_colormap(x)
    lut = sampling.plasma_lut()
    want = [0, 0, 0, 1, 128, 255, 255, 255, 0]
    for i, w in enumerate(want):
        assert tuple(rgb[0, i]) == tuple(lut[w]), (i, w)
    assert tuple(rgb[0, 9]) == (0, 0, 0)                                         # NaN -> "bad" colour
    assert rgb.dtype == np.uint8 and rgb.shape == (1, 10, 3)


def test_colour_uses_one_global_min_max_over_all_concepts():
    maps = np.zeros((3, 4, 4), dtype=np.float32)
    maps[0] += 1.0
    maps[1] += 3.0
    maps[2, 0, 0] = 5.0
    maps[2, 1, 1] = -1.0
    imgs = _colour(maps, "plasma")
    lut = sampling.plasma_lut()
    assert len(imgs) == 3 and imgs[0].size == (4, 4) and imgs[0].mode == "RGB"
    a = [np.asarray(im) for im in imgs]
    # global range [-1, 5]: concept 0 is flat at (1+1)/6, concept 1 flat at 4/6 — NOT stretched per concept
    assert (a[0] == lut[int(2 / 6 * 256)]).all() and (a[1] == lut[int(4 / 6 * 256)]).all()
    assert tuple(a[2][0, 0]) == tuple(lut[255]) and tuple(a[2][1, 1]) == tuple(lut[0])
