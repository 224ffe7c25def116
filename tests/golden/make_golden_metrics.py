"""Golden vectors for the segmentation metrics, from the REFERENCE's own functions
(concept_attention/utils.py:48-108: get_ap_scores, batch_pix_accuracy, batch_intersection_union).

Run in the build container only (needs /root/reference and scikit-learn, the reference's
dependency for average_precision_score):
    python tests/golden/make_golden_metrics.py
Writes tests/golden/metrics.npz; inputs are re-derived from seeds by tests/golden_inputs.py.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(os.path.dirname(HERE)))


def main():
    import make_golden
    make_golden.install_stubs()
    sys.path.insert(0, "/root/reference")
    from concept_attention.utils import batch_intersection_union, batch_pix_accuracy, get_ap_scores
    import sklearn
    import golden_inputs as gi

    out = dict(torch_version=torch.__version__, sklearn_version=sklearn.__version__,
               reference_commit="401dc36")
    cases = gi.metric_inputs()
    for name in ("binary", "voc"):
        pred, tgt, nclass = cases[name]
        correct, labeled = batch_pix_accuracy(pred, tgt)
        inter, union = batch_intersection_union(pred, tgt, nclass)
        out[name + "_correct"], out[name + "_labeled"] = np.int64(correct), np.int64(labeled)
        out[name + "_inter"], out[name + "_union"] = inter.astype(np.int64), union.astype(np.int64)
    scores, labels = cases["ap"]
    out["ap"] = np.asarray(get_ap_scores(scores, labels), dtype=np.float64)
    np.savez_compressed(os.path.join(HERE, "metrics.npz"), **out)
    print({k: (v if np.ndim(v) == 0 else np.asarray(v).tolist()) for k, v in out.items()})


if __name__ == "__main__":
    main()
