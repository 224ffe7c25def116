"""GPU parity: K6 segmentation counters and the device AP against the oracle and the
reference-generated golden vectors (tests/golden/metrics.npz).  Integer work: bit-exact."""
import os

import numpy as np
import pytest
import torch

import golden_inputs as gi
from oracle import metrics as om

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "metrics.npz")


@pytest.fixture(scope="module")
def metrics():
    assert torch.cuda.is_available(), "GPU tests need a B200"
    from conceptattention_b200 import metrics as m
    return m


def test_counts_match_reference_golden(metrics):
    g = np.load(GOLD)
    cases = gi.metric_inputs()
    for name in ("binary", "voc"):
        pred, tgt, nclass = cases[name]
        pred, tgt = pred.cuda(), tgt.cuda()
        assert metrics.batch_pix_accuracy(pred, tgt) == (int(g[name + "_correct"]), int(g[name + "_labeled"]))
        inter, union = metrics.batch_intersection_union(pred, tgt, nclass)
        assert np.array_equal(inter, g[name + "_inter"]) and np.array_equal(union, g[name + "_union"])
    scores, labels = cases["ap"]
    np.testing.assert_allclose(metrics.get_ap_scores(scores.cuda(), labels.cuda()), g["ap"], rtol=0, atol=1e-12)


@pytest.mark.parametrize("n,nclass", [(0, 3), (1, 1), (255, 2), (4097, 21), (1 << 20, 150), (300000, 1024),
                                      (300000, 1500), (9_000_000, 21)])
def test_counts_match_oracle(metrics, n, nclass):
    """Ragged sizes, one class, the shared-memory limit and the global-atomics form above it."""
    g = torch.Generator().manual_seed(n + nclass)
    tgt = torch.randint(-1, nclass, (n,), generator=g).float()
    pred = torch.randint(0, nclass + 2, (n,), generator=g).float()          # some labels out of range
    pred = torch.where(torch.rand(n, generator=g) > 0.4, tgt.clamp(min=0), pred)
    correct, labeled, inter, parea, larea = metrics.segmentation_counts(pred.cuda(), tgt.cuda(), nclass)
    assert (correct, labeled) == om.pix_accuracy(pred.numpy(), tgt.numpy())
    o_inter, o_union = om.intersection_union(pred.numpy(), tgt.numpy(), nclass)
    assert np.array_equal(inter, o_inter) and np.array_equal(parea + larea - inter, o_union)
    assert labeled == int((tgt >= 0).sum()) and int(larea.sum()) == labeled     # every labelled pixel binned


def test_all_ignored_and_bad_arguments(metrics):
    tgt = torch.full((64, 64), -1.0, device="cuda")
    correct, labeled, inter, parea, larea = metrics.segmentation_counts(torch.zeros_like(tgt), tgt, 4)
    assert (correct, labeled) == (0, 0) and not inter.any() and not parea.any() and not larea.any()
    with pytest.raises(ValueError):
        metrics.segmentation_counts(torch.zeros(4, device="cuda"), torch.zeros(5, device="cuda"), 2)
    with pytest.raises(RuntimeError):
        metrics.segmentation_counts(torch.zeros(4), torch.zeros(4), 2)          # no CPU path
    from conceptattention_b200 import _cabi
    assert _cabi.load().ca_segmentation_counts(0, 0, 8, 2, 0, 0) != 0           # null pointers refused


@pytest.mark.parametrize("n,levels", [(1, 2), (17, 2), (500, 4), (65536, 1000), (65536, 10 ** 9)])
def test_average_precision_matches_oracle(metrics, n, levels):
    rng = np.random.default_rng(n + levels)
    s = np.round(rng.random(n) * levels) / levels
    y = (rng.random(n) > 0.6).astype(np.float64)
    y[0] = 1.0
    got = metrics.average_precision(torch.from_numpy(y).cuda(), torch.from_numpy(s).cuda())
    assert abs(got - om.average_precision(y, s)) < 1e-12


def test_meter_matches_reference_loop(metrics):
    """SegmentationMeter against the running totals of run_experiment.py:206-234 restated with the oracle."""
    g = torch.Generator().manual_seed(8)
    meter = metrics.SegmentationMeter(nclass=2)
    correct = label = 0
    inter, union, aps = np.zeros(2), np.zeros(2), []
    for _ in range(3):
        coeff = torch.rand(224, 224, generator=g)
        mask = (coeff > coeff.mean()).float()
        tgt = (torch.rand(224, 224, generator=g) > 0.5).float()
        meter.update(mask.cuda(), tgt.cuda(), coeff.cuda())
        m2, t2 = torch.stack((1 - mask, mask)).numpy(), torch.stack((1 - tgt, tgt)).numpy()
        c, l = om.pix_accuracy(m2, t2)
        i, u = om.intersection_union(m2, t2, 2)
        correct, label, inter, union = correct + c, label + l, inter + i, union + u
        aps += om.ap_scores(torch.stack((1 - coeff, coeff)).unsqueeze(0).numpy(), tgt.unsqueeze(0).numpy())
    eps = np.spacing(1, dtype=np.float64)
    s = meter.summary()
    assert s["pixAcc"] == correct / (eps + label)
    assert s["mIoU"] == float((inter / (eps + union)).mean())
    assert abs(s["mAP"] - np.mean(aps)) < 1e-12
