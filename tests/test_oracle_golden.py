"""Pin the CPU oracle to outputs of the reference's own modules
(tests/golden/*.npz, produced by tests/golden/make_golden.py)."""
import os

import numpy as np
import pytest
import torch

import golden_inputs as gi
import oracle
from conceptattention_b200.weights import CONFIGS, init_state_dict

torch.set_grad_enabled(False)


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


def _close(a, b, rtol=2e-4, atol=2e-5):
    a = a.numpy() if isinstance(a, torch.Tensor) else a
    np.testing.assert_allclose(a, b, rtol=rtol, atol=atol)


@pytest.mark.parametrize("cfg", ["tiny", "tiny-dev"])
def test_dit_forward_matches_reference(golden_dir, cfg):
    g = _load(golden_dir, f"dit_{cfg}.npz")
    p = CONFIGS[cfg]
    sd = init_state_dict(p, seed=gi.WEIGHT_SEED, norm_scale_jitter=0.25)
    inp = gi.dit_inputs(p, gi.TINY_SIDE, gi.TINY_T, gi.TINY_C, gi.INPUT_SEED)
    t, gd = torch.full((1,), 0.75), torch.full((1,), 3.5)
    pred, vecs = oracle.dit_forward(sd, p.asdict(), inp["img"], inp["img_ids"], inp["txt"],
                                    inp["txt_ids"], inp["concepts"], inp["concept_ids"],
                                    inp["concept_vec"], t, inp["vec"], gd)
    _close(pred, g["pred"])
    for k in oracle.VECTOR_KEYS:
        _close(vecs[k], g[k])
    for name, kw in gi.ABLATIONS.items():
        none, v2 = oracle.dit_forward(sd, p.asdict(), inp["img"], inp["img_ids"], inp["txt"],
                                      inp["txt_ids"], inp["concepts"], inp["concept_ids"],
                                      inp["concept_vec"], t, inp["vec"], gd,
                                      stop_after_multimodal_attentions=True,
                                      joint_attention_kwargs=kw)
        assert none is None
        _close(v2["output_space_concept_vectors"], g["ablation_" + name])


def test_denoise_matches_reference(golden_dir):
    g = _load(golden_dir, "denoise_tiny.npz")
    p = CONFIGS["tiny"]
    sd = init_state_dict(p, seed=gi.WEIGHT_SEED, norm_scale_jitter=0.25)
    inp = gi.dit_inputs(p, gi.TINY_SIDE, gi.TINY_T, gi.TINY_C, gi.INPUT_SEED)
    ts = oracle.get_schedule(2, inp["img"].shape[1], shift=False)
    np.testing.assert_allclose(ts, g["schedule"], rtol=0, atol=0)
    x, st = oracle.denoise(sd, p.asdict(), inp["img"], inp["img_ids"], inp["txt"], inp["txt_ids"],
                           inp["vec"], ts, guidance=0.0, concepts=inp["concepts"],
                           concept_ids=inp["concept_ids"], concept_vec=inp["concept_vec"])
    _close(x, g["img"])
    _close(st["output_space_image_vectors"], g["out_img"])
    _close(st["output_space_concept_vectors"], g["out_con"])
    _close(st["cross_attention_image_vectors"], g["cross_img"])
    _close(st["cross_attention_concept_vectors"], g["cross_con"])


def test_heatmaps_match_reference(golden_dir):
    g = _load(golden_dir, "heatmaps.npz")
    for tag, (iv, cv) in gi.heatmap_inputs().items():
        h = oracle.heatmaps_from_vectors(iv, cv, gi.HM_LAYERS, gi.HM_TIMESTEPS, softmax=True)
        assert h.shape == (1, 5, 64, 64)
        _close(h, g[tag + "_softmax"], rtol=1e-5, atol=1e-6)
        h = oracle.heatmaps_from_vectors(iv, cv, gi.HM_LAYERS, gi.HM_TIMESTEPS, softmax=True,
                                         normalize_concepts=True)
        _close(h, g[tag + "_softmax_normcon"], rtol=1e-5, atol=1e-6)
    x = torch.randn(3, 5, 7, generator=torch.Generator().manual_seed(5))
    _close(oracle.linear_normalization(x, dim=-2), g["linear_normalization"], rtol=1e-6, atol=0)


def test_masks_schedule_noise_match_reference(golden_dir):
    g = _load(golden_dir, "masks.npz")
    coeff = gi.mask_inputs()
    assert np.array_equal(oracle.binary_masks(coeff).numpy(), g["masks_plain"])
    assert np.array_equal(oracle.binary_masks(coeff, apply_blur=True).numpy(), g["masks_blur"])
    am = oracle.argmax_mask(coeff)
    assert am.dtype == torch.int64 and np.array_equal(am.numpy(), g["argmax"])
    np.testing.assert_allclose(oracle.get_schedule(4, 1024, shift=True), g["schedule_shift"], rtol=1e-7)
    np.testing.assert_allclose(oracle.get_schedule(4, 1024, shift=False), g["schedule_noshift"], rtol=0)
    assert np.array_equal(oracle.get_noise(1, 256, 256, torch.float32, 7).numpy(), g["noise"])
    # add_noise_to_image (segmentation.py:85-113): x = t*noise + (1-t)*latent, bf16
    ts = oracle.get_schedule(4, 32 * 32 // 4, shift=True)
    t = ts[2]
    x = oracle.get_noise(1, 256, 256, torch.bfloat16, 3)
    noised = t * x + (1.0 - t) * (torch.zeros(1, 16, 32, 32) + 0.5).to(x.dtype)
    assert np.array_equal(noised.float().numpy(), g["noised"])
    np.testing.assert_allclose(ts[2:], g["noised_tail"], rtol=1e-7)


def test_unpinned_normalisers_properties():
    """sparsemax / entmax15 have no reference vectors (entmax absent): check the
    defining properties instead."""
    z = torch.randn(64, 7, generator=torch.Generator().manual_seed(1)) * 2
    for fn in (oracle.sparsemax, oracle.entmax15):
        p = fn(z, dim=-1)
        assert (p >= 0).all()
        np.testing.assert_allclose(p.sum(-1).numpy(), 1.0, atol=1e-5)
        assert (p.argmax(-1) == z.argmax(-1)).all()
    # sparsemax == Euclidean projection: optimality  z_i - p_i == tau on the support
    p = oracle.sparsemax(z, dim=-1)
    tau = ((z - p) * (p > 0)).sum(-1) / (p > 0).sum(-1)
    assert (((z - p) - tau[:, None]).abs() * (p > 0)).max() < 1e-5
    assert ((z - tau[:, None]) * (p == 0) <= 1e-6).all()
    # entmax15: p_i = max((z_i - tau)/2 ... )^2 form -> sqrt(p) is affine in z on the support
    p = oracle.entmax15(z, dim=-1)
    r = (z / 2 - torch.sqrt(p))
    tau = (r * (p > 0)).sum(-1) / (p > 0).sum(-1)
    assert ((r - tau[:, None]).abs() * (p > 0)).max() < 1e-5


def test_segmentation_metrics_match_reference(golden_dir):
    """oracle/metrics.py against the reference's utils.py:48-108 outputs (make_golden_metrics.py)."""
    from oracle import metrics as om
    g = np.load(os.path.join(golden_dir, "metrics.npz"))
    cases = gi.metric_inputs()
    for name in ("binary", "voc"):
        pred, tgt, nclass = cases[name]
        assert om.pix_accuracy(pred.numpy(), tgt.numpy()) == (int(g[name + "_correct"]), int(g[name + "_labeled"]))
        inter, union = om.intersection_union(pred.numpy(), tgt.numpy(), nclass)
        assert np.array_equal(inter, g[name + "_inter"]) and np.array_equal(union, g[name + "_union"])
    scores, labels = cases["ap"]
    np.testing.assert_allclose(om.ap_scores(scores.numpy(), labels.numpy()), g["ap"], rtol=0, atol=1e-12)


def test_average_precision_matches_sklearn_on_edge_cases():
    """The AP restatement against scikit-learn itself (the reference's dependency, utils.py:66)."""
    skm = pytest.importorskip("sklearn.metrics")
    from oracle import metrics as om
    rng = np.random.default_rng(3)
    for n, levels in ((1, 2), (17, 2), (500, 4), (4096, 1000), (4096, 10 ** 9)):
        s = np.round(rng.random(n) * levels) / levels
        y = (rng.random(n) > 0.6).astype(np.float64)
        y[0] = 1.0
        assert abs(om.average_precision(y, s) - skm.average_precision_score(y, s)) < 1e-12, (n, levels)
    s = rng.random(64)
    assert abs(om.average_precision(np.ones(64), s) - 1.0) < 1e-12      # all positive
