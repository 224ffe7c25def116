"""oracle/_ref — the reference's own modules placed by oracle/make_ref.py — reproduce the committed
golden vectors (made by tests/golden/make_golden.py from /root/reference) bit for bit, and agree with
the oracle port.  This is what lets the GPU box, where /root/reference does not exist, use them as the
comparator of the end-to-end parity tests and as bench.py's reference arms."""
import os

import numpy as np
import pytest
import torch

import golden_inputs as gi
import oracle
from oracle import reference_modules as rm
from conceptattention_b200.weights import CONFIGS, init_state_dict

pytestmark = pytest.mark.skipif(not rm.available(), reason="no oracle/_ref and no /root/reference")


def test_reference_modules_reproduce_golden_and_match_port(golden_dir):
    torch.set_grad_enabled(False)
    ns = rm.load()
    g = np.load(os.path.join(golden_dir, "dit_tiny.npz"))
    p = CONFIGS["tiny"]
    sd = init_state_dict(p, seed=gi.WEIGHT_SEED, norm_scale_jitter=0.25)
    model = rm.build_dit(p.asdict(), sd, "cpu", torch.float32)
    inp = gi.dit_inputs(p, gi.TINY_SIDE, gi.TINY_T, gi.TINY_C, gi.INPUT_SEED)
    t, gd = torch.full((1,), 0.75), torch.full((1,), 3.5)
    pred, vecs = model(img=inp["img"], img_ids=inp["img_ids"], txt=inp["txt"], txt_ids=inp["txt_ids"],
                       concepts=inp["concepts"], concept_ids=inp["concept_ids"], concept_vec=inp["concept_vec"],
                       y=inp["vec"], timesteps=t, guidance=gd)
    if str(g["torch_version"]) == torch.__version__:
        np.testing.assert_array_equal(pred.numpy(), g["pred"])
    np.testing.assert_allclose(pred.numpy(), g["pred"], rtol=1e-4, atol=1e-5)
    p2, v2 = oracle.dit_forward(sd, p.asdict(), inp["img"], inp["img_ids"], inp["txt"], inp["txt_ids"],
                                inp["concepts"], inp["concept_ids"], inp["concept_vec"], t, inp["vec"], gd)
    torch.testing.assert_close(p2, pred, rtol=1e-5, atol=1e-6)
    for k in oracle.VECTOR_KEYS:
        torch.testing.assert_close(v2[k], vecs[k], rtol=1e-5, atol=1e-6)


def test_heatmap_patch_only_changes_the_grid_size():
    """Patch (1) of oracle/reference_modules.py: h = w = isqrt(P) instead of 64; at P = 4096 the
    patched function IS the unmodified one (golden heatmaps.npz), at other sizes it matches the port."""
    torch.set_grad_enabled(False)
    ns = rm.load()
    gen = torch.Generator().manual_seed(5)
    iv, cv = torch.randn(2, 3, 1, 144, 32, generator=gen), torch.randn(2, 3, 1, 5, 32, generator=gen)
    got = ns.compute_heatmaps_from_vectors(iv, cv, layer_indices=[1, 2], timesteps=[0, 1])
    want = oracle.heatmaps_from_vectors(iv, cv, [1, 2], [0, 1])
    assert got.shape == (1, 5, 12, 12)
    torch.testing.assert_close(got, want, rtol=1e-6, atol=1e-7)


def test_make_ref_lists_the_import_closure():
    """Every reference file imported on the hot path is in oracle/make_ref.py's list (so the GPU box has it)."""
    import sys
    from oracle import make_ref
    rm.load()
    root = rm.root()
    loaded = {os.path.relpath(m.__file__, root) for m in list(sys.modules.values())
              if getattr(m, "__file__", None) and m.__file__.startswith(root + os.sep)}
    assert loaded <= set(make_ref.FILES), loaded - set(make_ref.FILES)
