"""Host logic of K2's stream-K tail on CPU (no GPU call): the planner and the range / piece
arithmetic that kernel and host share (`plan_streamk_core`, `sk_range`, `sk_piece` in
csrc/k2_attention.cu, exported through two debug entry points that make no CUDA call).

Properties: a plan keeps every main unit exactly once (whole or tail); whole units + concept units fill
whole waves (or everything is tail when there is less than one wave); the pieces of the tail cover every key step of every tail unit exactly once, contiguously,
in range order; a range touches at most two units; range lengths differ by at most one step."""
import ctypes as C

import pytest
from hypothesis import given, settings, strategies as st

from conceptattention_b200 import _cabi


@pytest.fixture(scope="module")
def lib():
    return _cabi.load()


def plan(lib, W, n_concept, n_steps, sms, auto):
    out = (C.c_int * 3)()
    lib.ca_debug_plan_streamk(W, n_concept, n_steps, sms, int(auto), out)
    return tuple(out)


def pieces(lib, n_steps, sk_units, sk_sms):
    out = (C.c_int * 3)()
    res = []
    for i in range(sk_sms):
        for k in (0, 1):
            if lib.ca_debug_streamk_piece(n_steps, sk_units, sk_sms, i, k, out):
                res.append((i, k, out[0], out[1], out[2]))
    return res


def test_measured_shapes(lib):
    # config 2: 408 pair units + 24 concept units, 34 key tiles: tail wave 0.76 full -> not split
    assert plan(lib, 408, 24, 34, 148, True) == (408, 0, 0)
    # config 5 on 8 GPUs: 195 pair units + 3 concept units, 130 key tiles: 145 whole + 50 split over 148
    assert plan(lib, 195, 3, 130, 148, True) == (145, 50, 148)
    # same shape without the measured thresholds (test mode) and with a tail more than half full
    assert plan(lib, 260, 0, 130, 148, True) == (260, 0, 0)
    assert plan(lib, 260, 0, 130, 148, False) == (148, 112, 148)
    # nothing to do: exact waves, fewer steps than ranges, degenerate inputs
    assert plan(lib, 296, 0, 130, 148, False) == (296, 0, 0)
    assert plan(lib, 10, 0, 2, 148, False) == (10, 0, 0)
    assert plan(lib, 0, 0, 130, 148, True) == (0, 0, 0)
    assert plan(lib, 5, 0, 130, 0, True) == (5, 0, 0)


@settings(max_examples=300, deadline=None)
@given(W=st.integers(1, 3000), n_concept=st.integers(0, 400), n_steps=st.integers(1, 200),
       sms=st.integers(1, 200), auto=st.booleans())
def test_plan_and_pieces_partition_the_tail(W, n_concept, n_steps, sms, auto):
    lib = _cabi.load()
    n_dp, sk_units, sk_sms = plan(lib, W, n_concept, n_steps, sms, auto)
    if sk_sms == 0:
        assert (n_dp, sk_units) == (W, 0)
        return
    assert sk_sms == sms and n_dp >= 0 and 0 < sk_units < sms and n_dp + sk_units == W
    whole = ((W + n_concept) // sms) * sms - n_concept        # whole units + concept units = whole waves
    assert n_dp == max(whole, 0)                              # (less than one wave in all: everything is tail)
    assert sk_units * n_steps // sms >= 4                     # never ranges shorter than 4 key steps
    if auto:
        assert 2 * sk_units <= sms and (sms - sk_units) * n_steps // sms >= 24
    covered = {u: [] for u in range(sk_units)}
    per_range = {}
    for i, k, unit, j_begin, n_local in pieces(lib, n_steps, sk_units, sk_sms):
        assert 0 <= unit < sk_units and n_local > 0 and 0 <= j_begin and j_begin + n_local <= n_steps
        covered[unit].append((j_begin, n_local, i, k))
        per_range.setdefault(i, []).append((k, unit, n_local))
    for unit, segs in covered.items():                        # every key step exactly once, in range order
        segs.sort()
        pos = 0
        for j_begin, n_local, _, _ in segs:
            assert j_begin == pos
            pos += n_local
        assert pos == n_steps
        assert [s[2] for s in segs] == sorted(s[2] for s in segs)
    lengths = []
    for i, ps in per_range.items():
        assert len(ps) <= 2
        if len(ps) == 2:
            assert ps[0][1] + 1 == ps[1][1]                   # piece 1 continues into the NEXT unit
        lengths.append(sum(p[2] for p in ps))
    assert max(lengths) - min(lengths) <= 1 and sum(lengths) == sk_units * n_steps
