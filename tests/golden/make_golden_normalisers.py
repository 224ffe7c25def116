"""Closed-form sparsemax / entmax-1.5 vectors for C = 2 and C = 3 (plain Python float64, no torch,
no oracle): the `entmax` package the reference imports (concept_attention_pipeline.py:5,66-69) is
neither vendored nor installed here, so these two normalisers are pinned against the DEFINITIONS
instead — sparsemax(z) = argmin_p ||p - z||^2 on the simplex (Martins & Astudillo 2016),
entmax15(z) = [z/2 - tau]_+^2 with sum 1 (Peters et al. 2019) — solved by hand per support size:

  sparsemax, support k (sorted z1 >= z2 >= ...): tau = (z1 + .. + zk - 1) / k, valid iff zk > tau
  entmax15,  support k, x = z / 2, m = mean(x1..xk), ss = sum (xi - m)^2:
             tau = m - sqrt((1 - ss) / k), valid iff xk > tau            (k = 1: one-hot)

usage: python tests/golden/make_golden_normalisers.py  ->  tests/golden/normalisers.npz"""
import math
import os

import numpy as np

CASES = [
    # C = 2: both in the support / exactly at the boundary / one-hot / tie / large common offset
    [0.3, -0.2], [1.0, 0.0], [2.5, 0.1], [0.7, 0.7], [101.25, 100.5], [-3.0, -2.1], [0.0, 1.999],
    # C = 3: support 3, 2 and 1; ties; boundary of the support; negative / large entries
    [0.1, 0.0, -0.1], [0.5, 0.3, -0.4], [1.2, 1.1, -0.5], [3.0, 0.2, 0.1], [0.4, 0.4, 0.4],
    [1.0, 1.0, -1.0], [2.0, 0.0, 0.0], [-5.0, -5.5, -7.0], [50.2, 49.9, 49.4], [0.9, 0.1, -0.1],
]


def sparsemax(z):
    order = sorted(range(len(z)), key=lambda i: -z[i])
    zs = [z[i] for i in order]
    for k in range(len(z), 0, -1):
        tau = (sum(zs[:k]) - 1.0) / k
        if zs[k - 1] > tau:
            break
    return [max(v - tau, 0.0) for v in z]


def entmax15(z):
    x = [v / 2.0 for v in z]
    xs = sorted(x, reverse=True)
    for k in range(len(z), 0, -1):
        m = sum(xs[:k]) / k
        ss = sum((v - m) ** 2 for v in xs[:k])
        if ss > 1.0:
            continue
        tau = m - math.sqrt((1.0 - ss) / k)
        if xs[k - 1] > tau:
            break
    return [max(v - tau, 0.0) ** 2 for v in x]


def main():
    out = {}
    for c in (2, 3):
        zs = [z for z in CASES if len(z) == c]
        out[f"z{c}"] = np.asarray(zs, dtype=np.float64)
        out[f"sparsemax{c}"] = np.asarray([sparsemax(z) for z in zs], dtype=np.float64)
        out[f"entmax15_{c}"] = np.asarray([entmax15(z) for z in zs], dtype=np.float64)
        for name in (f"sparsemax{c}", f"entmax15_{c}"):
            assert np.allclose(out[name].sum(1), 1.0, atol=1e-12), name
    np.savez(os.path.join(os.path.dirname(os.path.abspath(__file__)), "normalisers.npz"), **out)
    for k, v in out.items():
        print(k, v.round(6).tolist())


if __name__ == "__main__":
    main()
