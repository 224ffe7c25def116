"""Oracle: segmentation metrics (TEST INFRASTRUCTURE ONLY, see oracle/__init__.py).

numpy restatement of concept_attention/utils.py:48-108.  `average_precision` restates
scikit-learn's average_precision_score for binary targets (sklearn/metrics/_ranking.py:
_binary_clf_curve + precision_recall_curve + AP = sum_k (R_k - R_{k-1}) P_k), the third-party
routine the reference calls (utils.py:66); pinned against tests/golden/metrics.npz, which was
produced by the reference's get_ap_scores with scikit-learn 1.9.0.
"""
from __future__ import annotations

import numpy as np


def pix_accuracy(predict, target):
    """utils.py:71-85 -> (pixel_correct, pixel_labeled)."""
    p, t = np.asarray(predict) + 1, np.asarray(target) + 1
    return int(np.sum((p == t) * (t > 0))), int(np.sum(t > 0))


def _hist(v, nclass):
    """np.histogram(bins=nclass, range=(1, nclass)) for integer-valued v: bin v-1, the rest dropped."""
    v = np.asarray(v).reshape(-1)
    keep = (v >= 1) & (v <= nclass)
    return np.bincount(v[keep].astype(np.int64) - 1, minlength=nclass).astype(np.int64)


def intersection_union(predict, target, nclass):
    """utils.py:88-108 -> (area_inter[nclass], area_union[nclass])."""
    p, t = np.asarray(predict) + 1, np.asarray(target) + 1
    p = p * (t > 0).astype(p.dtype)
    inter = p * (p == t)
    a_i, a_p, a_l = _hist(inter, nclass), _hist(p, nclass), _hist(t, nclass)
    return a_i, a_p + a_l - a_i


def average_precision(y_true, y_score):
    """Binary AP with tied scores grouped into one threshold (stable descending sort)."""
    y_true, y_score = np.asarray(y_true, np.float64), np.asarray(y_score, np.float64)
    order = np.argsort(-y_score, kind="mergesort")
    y_true, y_score = y_true[order], y_score[order]
    last = np.r_[np.nonzero(np.diff(y_score))[0], y_true.size - 1]   # last index of each tie group
    tps = np.cumsum(y_true)[last]
    precision = tps / (last + 1.0)
    recall = tps / tps[-1] if tps[-1] > 0 else np.ones_like(tps)
    return float(np.sum(np.diff(np.r_[0.0, recall]) * precision))


def ap_scores(predict, target, ignore_index=-1):
    """utils.py:48-69: predict [n,K,h,w] scores, target [n,h,w] labels -> list of n APs."""
    total = []
    for pred, tgt in zip(np.asarray(predict), np.asarray(target)):
        expand = np.broadcast_to(tgt[None], pred.shape).reshape(-1)
        onehot = (np.arange(pred.shape[0])[:, None, None] == np.clip(tgt, 0, None)[None]).astype(np.float64)
        keep = expand != ignore_index
        total.append(float(np.nan_to_num(
            average_precision(onehot.reshape(-1)[keep], np.nan_to_num(pred.reshape(-1))[keep]))))
    return total
