"""CPU-side checks of the drop-in boundary: the C-ABI library builds (nvcc cross-compiles
without a GPU), loads, and exports every symbol include/conceptattention_b200.h declares; the
Python binding covers the same set; the product fails loudly (no CPU / eager fallback) when
the library is missing or a host tensor is passed.  No compute call is made here."""
import ctypes
import os
import re

import pytest
import torch

from conceptattention_b200 import _cabi, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "conceptattention_b200.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ca_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    path = build.build()
    assert os.path.exists(path)
    return ctypes.CDLL(path)


def test_header_declares_the_expected_surface():
    names = _declared()
    for must in ("ca_ln_modulate", "ca_gate_residual_ln_modulate", "ca_qkv_prep", "ca_joint_attention",
                 "ca_heatmap_accumulate", "ca_heatmap_logits", "ca_norm_accumulate",
                 "ca_concept_linear_normalize", "ca_finalize_heatmaps", "ca_gaussian_blur3",
                 "ca_argmax_mask", "ca_binary_masks", "ca_heatmaps_from_vectors_host",
                 "ca_device_check", "ca_status_string", "ca_version", "ca_last_cuda_error"):
        assert must in names, must


def test_library_exports_every_declared_symbol(lib):
    missing = [n for n in _declared() if not hasattr(lib, n)]
    assert not missing, missing


def test_binding_covers_every_declared_symbol():
    assert sorted(_cabi.SIGNATURES) == _declared()
    h = _cabi.load()                       # binds argtypes for all of them
    assert h.ca_version().decode().startswith("conceptattention_b200")


def test_status_strings(lib):
    lib.ca_status_string.restype = ctypes.c_char_p
    lib.ca_status_string.argtypes = [ctypes.c_int]
    seen = {lib.ca_status_string(code).decode() for code in (0, -1, -2, -3, -4, -5)}
    assert len(seen) == 6 and "ok" in seen


def test_no_fallback_when_library_is_missing(monkeypatch, tmp_path):
    monkeypatch.setattr(_cabi, "_lib", None)
    monkeypatch.setattr(_cabi, "_LIB_PATH", str(tmp_path / "libconceptattention_b200.so"))
    with pytest.raises(_cabi.ConceptAttentionLibraryError):
        _cabi.load()


def test_ops_reject_host_tensors():
    from conceptattention_b200 import ops
    x = torch.zeros(1, 4, 256, dtype=torch.bfloat16)
    with pytest.raises(ValueError):
        ops.ln_modulate(x, torch.zeros(1, 1536, dtype=torch.bfloat16))
    with pytest.raises(ValueError):
        ops.joint_attention(*(torch.zeros(1, 2, 8, 128, dtype=torch.bfloat16),) * 3, 4, 2, 2)
    with pytest.raises(ValueError):
        ops.argmax_mask(torch.zeros(1, 2, 4, 4))


def test_pipeline_argument_validation_mirrors_reference():
    """Unknown attention_norm raises ValueError like concept_attention_pipeline.py:70-71."""
    from conceptattention_b200.dit import HeatmapAccumulator
    with pytest.raises(ValueError):
        HeatmapAccumulator(1, 2, 16, "cpu", [0], None, softmax=False, attention_norm="bogus")
