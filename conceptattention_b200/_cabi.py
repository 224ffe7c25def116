"""ctypes binding of libconceptattention_b200.so (include/conceptattention_b200.h).

The library is the product: if it cannot be loaded this module raises — there is
no eager/PyTorch fallback anywhere in the package.
"""
from __future__ import annotations

import ctypes as C
import os

import torch  # noqa: F401  (loads libcudart.so.12 that the library links against)

_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libconceptattention_b200.so")

CA_OK = 0
NORM = {"softmax": 0, "sparsemax": 1, "entmax15": 2, "none": 3}
ATTN_BOTH, ATTN_CROSS_ONLY, ATTN_SELF_ONLY, ATTN_NEITHER = 0, 1, 2, 3
K2_AUTO, K2_V3, K2_V7, K2_V11, K2_V11P = 0, 1, 2, 3, 4
MAX_CONCEPTS = 32            # CA_MAX_CONCEPTS

_p, _i, _i64, _f = C.c_void_p, C.c_int, C.c_int64, C.c_float

# name -> argtypes; restype is int (ca_status) unless listed in _RESTYPES
SIGNATURES = {
    "ca_version": [],
    "ca_status_string": [_i],
    "ca_last_cuda_error": [],
    "ca_device_check": [_i],
    "ca_ln_modulate": [_p, _p, _i, _i, _i, _p, _p, _i, _i64, _i, _i, _p],
    "ca_gate_residual_ln_modulate": [_p, _p, _i64, _p, _i, _i, _i, _p, _p, _i, _i64, _i, _i, _i, _p],
    "ca_qkv_prep": [_p, _i64, _i, _p, _i64, _i, _p, _p, _p, _p, _p, _i, _i, _p, _p, _p, _p, _p],
    "ca_joint_attention": [_p, _p, _p, _i, _i, _i, _i, _i, _i, _p, _p, _p],
    "ca_joint_attention_workspace_bytes": [_i] * 6,
    "ca_joint_attention_ws": [_p, _p, _p, _i, _i, _i, _i, _i, _i, _p, _p, _p, C.c_size_t, _p],
    "ca_joint_attention_workspace_bytes_ex": [_i] * 8,
    "ca_joint_attention_ex": [_p, _p, _p, _i, _i, _i, _i, _i, _i, _p, _p, _p, C.c_size_t, _i, _i, _p],
    "ca_joint_attention_noalloc": [_p, _p, _p, _i, _i, _i, _i, _i, _i, _p, _p, _i, _i, _p],
    "ca_heatmap_accumulate": [_p, _i64, _i64, _i64, _p, _i64, _i64, _i64, _i, _i, _i, _i, _i, _i, _i, _p, _p],
    "ca_heatmap_accumulate_pair": [_p, _i64, _i64, _i64, _p, _i64, _i64, _i64, _i, _i, _i, _i, _i, _i, _i, _p, _p,
                                   _i64, _p, _i, _p],
    "ca_heatmap_accumulate_both": [_p, _i64, _i64, _i64, _p, _i64, _i64, _i64, _i, _i, _p, _i, _p, _i64, _i64, _i64, _p, _i64, _i64, _i64, _i, _i, _p, _i, _i, _i, _i, _i, _i, _p, _i64, _p, _p],
    "ca_norm_accumulate_pair": [_p, _i, _i, _i, _i, _i, _p, _p, _p],
    "ca_heatmap_logits": [_p, _i64, _i64, _i64, _p, _i64, _i64, _i64, _i, _i, _i, _i, _i, _p, _p],
    "ca_norm_accumulate": [_p, _i, _i, _i, _i, _i, _p, _p],
    "ca_concept_linear_normalize": [_p, _p, _i, _i64, _p],
    "ca_finalize_heatmaps": [_p, _p, _i64, _f, _i, _p],
    "ca_gaussian_blur3": [_p, _p, _i, _i, _i, _p],
    "ca_argmax_mask": [_p, _p, _i, _i, _i, _i, _p],
    "ca_binary_masks": [_p, _p, _i, _i, _p],
    "ca_segmentation_counts": [_p, _p, _i64, _i, _p, _p],
    "ca_qkv_prep_grouped": [_p, _i64, _i, _p, _i64, _i, _p, _p, _p, _p, _p, _i, _i, _p, _p, _p, _p, _i, _i64, _p],
    "ca_seq_unpack_qkv": [_p, _i, _i, _i, _i, _i, _i, _p, _p, _p, _p],
    "ca_seq_pack_out": [_p, _p, _i, _i, _i, _i, _i, _p, _p],
    "ca_seq_unpack_out": [_p, _i, _i, _i, _i, _i, _p, _p, _p],
    "ca_seq_gather_qkv": [_p, C.c_int64, _i, _i, _i, _i, _i, _i, _p, _p, _p, _p],
    "ca_seq_gather_out": [_p, C.c_int64, C.c_int64, _i, _i, _i, _i, _i, _i, _p, _p, _p],
    "ca_heatmaps_from_vectors_host": [_p, _p, _i, _i, _i, _i, _i, _i, _i, _i, _p, _i, _p, _i, _i, _i, _p],
}
_RESTYPES = {"ca_version": C.c_char_p, "ca_status_string": C.c_char_p,
             "ca_joint_attention_workspace_bytes": C.c_size_t,
             "ca_joint_attention_workspace_bytes_ex": C.c_size_t}

_lib = None


class ConceptAttentionLibraryError(RuntimeError):
    pass


def lib_path() -> str:
    return _LIB_PATH


def load():
    """dlopen the library and bind every declared symbol (raises if absent)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        raise ConceptAttentionLibraryError(
            f"{_LIB_PATH} not found: run `python -m conceptattention_b200.build` "
            "(there is no CPU / PyTorch fallback)")
    lib = C.CDLL(_LIB_PATH)
    for name, argtypes in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the .so is stale
        fn.argtypes = argtypes
        fn.restype = _RESTYPES.get(name, C.c_int)
    lib.ca_debug_plan_streamk.argtypes = [_i, _i, _i, _i, _i, _p]
    lib.ca_debug_plan_streamk.restype = None
    lib.ca_debug_streamk_piece.argtypes = [_i, _i, _i, _i, _i, _p]
    lib.ca_debug_streamk_piece.restype = _i
    _lib = lib
    return lib


def check(status: int, what: str):
    if status != CA_OK:
        lib = load()
        msg = lib.ca_status_string(status).decode()
        extra = f" (cuda error {lib.ca_last_cuda_error()})" if status == -5 else ""
        raise ConceptAttentionLibraryError(f"{what}: {msg}{extra}")
