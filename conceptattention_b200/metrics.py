"""Segmentation metrics on the device.

Same names, arguments and return values as concept_attention/utils.py:48-108
(`get_ap_scores`, `batch_pix_accuracy`, `batch_intersection_union`) and the running totals of
experiments/imagenet_segmentation/run_experiment.py:134-234 (`SegmentationMeter`).  The reference
copies predictions and labels to the host and counts with numpy; here one K6 launch
(`ca_segmentation_counts`) produces every counter from the device-resident label maps and only the
2 + 3*nclass integers cross PCIe.  The ranking step of average precision is a device sort + prefix
sum (library ops, like the GEMMs).  No CPU path: without the CUDA library these raise.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _cabi
from .ops import check, _stream


def _as_f32(x, device=None):
    x = torch.as_tensor(x)
    if device is not None:
        x = x.to(device)
    if not x.is_cuda:
        raise RuntimeError("segmentation metrics run on the GPU: pass CUDA tensors (no CPU path)")
    return x.to(torch.float32).contiguous()


def segmentation_counts(predict, target, nclass):
    """-> int64 numpy (pixel_correct, pixel_labeled, area_inter[nclass], area_pred[nclass], area_lab[nclass])."""
    target = _as_f32(target)
    predict = _as_f32(predict, target.device)
    if predict.shape != target.shape:
        raise ValueError(f"predict {tuple(predict.shape)} and target {tuple(target.shape)} differ")
    counts = torch.empty(2 + 3 * nclass, device=target.device, dtype=torch.int64)
    check(_cabi.load().ca_segmentation_counts(predict.data_ptr(), target.data_ptr(), predict.numel(),
                                              int(nclass), counts.data_ptr(), _stream(target)),
          "ca_segmentation_counts")
    c = counts.cpu().numpy()
    return int(c[0]), int(c[1]), c[2:2 + nclass], c[2 + nclass:2 + 2 * nclass], c[2 + 2 * nclass:]


def batch_pix_accuracy(predict, target):
    """utils.py:71-85 -> (pixel_correct, pixel_labeled)."""
    correct, labeled, *_ = segmentation_counts(predict, target, 1)
    return correct, labeled


def batch_intersection_union(predict, target, nclass):
    """utils.py:88-108 -> (area_inter, area_union), int64 arrays of length nclass."""
    _, _, inter, pred, lab = segmentation_counts(predict, target, nclass)
    return inter, pred + lab - inter


def average_precision(y_true, y_score):
    """Binary AP (what sklearn's average_precision_score computes, utils.py:66): stable descending
    sort, tied scores share one threshold, AP = sum_k (R_k - R_{k-1}) P_k in fp64."""
    y_score = torch.nan_to_num(y_score.to(torch.float64).reshape(-1))
    y_true = y_true.to(torch.float64).reshape(-1)
    n = y_score.numel()
    if n == 0:
        return 0.0
    s, order = torch.sort(y_score, descending=True, stable=True)
    tps_all = torch.cumsum(y_true[order], 0)
    last = torch.ones(n, dtype=torch.bool, device=s.device)
    last[:-1] = s[1:] != s[:-1]
    idx = torch.nonzero(last).squeeze(1)
    tps = tps_all[idx]
    precision = tps / (idx + 1).to(torch.float64)
    recall = tps / tps[-1] if tps[-1] > 0 else torch.ones_like(tps)
    prev = torch.cat([recall.new_zeros(1), recall[:-1]])
    return float(torch.nan_to_num(((recall - prev) * precision).sum()).item())


def get_ap_scores(predict, target, ignore_index=-1):
    """utils.py:48-69: predict [n,K,h,w] scores, target [n,h,w] labels -> list of n APs."""
    target = torch.as_tensor(target)
    predict = torch.as_tensor(predict).to(target.device)
    if not target.is_cuda:
        raise RuntimeError("segmentation metrics run on the GPU: pass CUDA tensors (no CPU path)")
    total = []
    for pred, tgt in zip(predict, target):
        K = pred.shape[0]
        onehot = torch.arange(K, device=tgt.device).view(K, *([1] * tgt.dim())) == tgt.clamp(min=0).long()[None]
        keep = (tgt != ignore_index)[None].expand_as(pred)
        total.append(average_precision(onehot[keep], pred[keep]))
    return total


class SegmentationMeter:
    """Running pixAcc / mIoU / mAP of run_experiment.py:134-234 (same epsilon, same means)."""

    def __init__(self, nclass=2):
        self.nclass = nclass
        self.correct = self.label = 0
        self.inter = np.zeros(nclass, np.float64)
        self.union = np.zeros(nclass, np.float64)
        self.ap = []

    def update(self, mask, target, coefficients=None):
        """mask/target: binary maps [h,w]; coefficients: scores [h,w] in [0,1] (optional -> AP).
        Stacks (1 - x, x) as the reference does before counting (run_experiment.py:206-217)."""
        target = _as_f32(target)
        mask = _as_f32(mask, target.device)
        correct, labeled, inter, pred, lab = segmentation_counts(
            torch.stack((1 - mask, mask)), torch.stack((1 - target, target)), self.nclass)
        self.correct += correct
        self.label += labeled
        self.inter += inter
        self.union += pred + lab - inter
        if coefficients is not None:
            c = _as_f32(coefficients, target.device)
            self.ap += get_ap_scores(torch.stack((1 - c, c)).unsqueeze(0), target.unsqueeze(0))
        return self

    def summary(self):
        eps = np.spacing(1, dtype=np.float64)
        iou = self.inter / (eps + self.union)
        return {"pixAcc": float(self.correct / (eps + self.label)), "mIoU": float(iou.mean()),
                "mAP": float(np.mean(self.ap)) if self.ap else float("nan")}
