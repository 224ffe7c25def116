"""GPU parity: heatmap kernels (K3/K4), masks (K5), the host-buffer C-ABI call and
the whole engine against the oracle and the reference-generated golden vectors."""
import ctypes
import math
import os

import numpy as np
import pytest
import torch

import golden_inputs as gi
import oracle
from conceptattention_b200.weights import CONFIGS, init_state_dict

pytestmark = pytest.mark.gpu
BF16 = torch.bfloat16


@pytest.fixture(scope="module")
def ops():
    assert torch.cuda.is_available(), "GPU tests need a B200"
    from conceptattention_b200 import ops as _ops
    return _ops


def rnd(*shape, seed=0, scale=1.0, dtype=BF16):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(dtype)


def _oracle_maps(iv, cv, norm, per_head):
    """one (t,l) pair through the oracle; iv [B,P,K] or [B,H,P,d]."""
    iv, cv = iv.float()[None, None], cv.float()[None, None]
    if norm == "none":
        if per_head:
            return torch.einsum("bhpd,bhcd->bcp", iv[0, 0], cv[0, 0])
        return torch.einsum("bpd,bcd->bcp", iv[0, 0], cv[0, 0])
    h = oracle.heatmaps_from_vectors(iv, cv, [0], [0], softmax=(norm == "softmax"),
                                     attention_norm=norm)
    return h.reshape(h.shape[0], h.shape[1], -1)


@pytest.mark.parametrize("norm", ["softmax", "sparsemax", "entmax15", "none"])
@pytest.mark.parametrize("B,P,C,K", [(1, 144, 3, 256), (2, 1000, 4, 3072), (1, 333, 8, 512),
                                     (1, 257, 20, 1024), (1, 130, 32, 3072), (1, 64, 1, 128)])
def test_heatmap_accumulate(ops, norm, B, P, C, K):
    iv, cv = rnd(B, P, K, seed=1), rnd(B, C, K, seed=2, scale=3.0 / math.sqrt(K))
    acc = torch.zeros(B, C, P, device="cuda")
    ops.heatmap_accumulate(iv.cuda(), cv.cuda(), acc, norm)
    ops.heatmap_accumulate(iv.cuda(), cv.cuda(), acc, norm)          # accumulates
    ref = 2 * _oracle_maps(iv, cv, norm, False)
    torch.testing.assert_close(acc.cpu(), ref, rtol=2e-4, atol=2e-4)


@pytest.mark.parametrize("B,P,C,K", [(1, 4096, 4, 3072), (2, 1000, 20, 3072), (1, 300, 32, 1024),
                                     (3, 129, 16, 128), (1, 16384, 32, 3072)])
def test_heatmap_tc_matches_simt_and_oracle(ops, B, P, C, K):
    """The TMA/tcgen05 streaming kernel (cluster split-K, DSMEM reduce; vector length a multiple
    of 64) and the SIMT kernel (every other length: here the same vectors zero-padded by 8
    columns, which changes no dot product) compute the same maps; both match the oracle."""
    iv, cv = rnd(B, P, K, seed=11), rnd(B, C, K, seed=12, scale=3.0 / math.sqrt(K))
    pad = lambda t: torch.cat((t, torch.zeros(*t.shape[:-1], 8, dtype=t.dtype)), -1)
    out = []
    for a, b in ((iv, cv), (pad(iv), pad(cv))):
        acc = torch.zeros(B, C, P, device="cuda")
        ops.heatmap_accumulate(a.cuda(), b.cuda(), acc, "softmax")
        logits = ops.heatmap_logits(a.cuda(), b.cuda())
        out.append((acc.cpu(), logits.cpu()))
    torch.testing.assert_close(out[0][0], out[1][0], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(out[0][1], out[1][1], rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(out[0][0], _oracle_maps(iv, cv, "softmax", False), rtol=2e-4, atol=2e-4)


@pytest.mark.parametrize("norm", ["softmax", "sparsemax"])
def test_heatmap_accumulate_per_head_strided(ops, norm):
    """cross-attention space: per-head vectors read in place from a [B,H,N,128] buffer."""
    B, H, P, T, C = 2, 3, 200, 40, 5
    qpre = rnd(B, H, P + T + C, 128, seed=3, scale=0.3)
    g = qpre.cuda()
    acc = torch.zeros(B, C, P, device="cuda")
    ops.heatmap_accumulate(g[:, :, :P], g[:, :, P + T:], acc, norm, per_head=True)
    ref = _oracle_maps(qpre[:, :, :P], qpre[:, :, P + T:], norm, True)
    torch.testing.assert_close(acc.cpu(), ref, rtol=2e-4, atol=2e-4)


def test_logits_then_norm_equals_fused(ops):
    """head-sharded form: per-slice partial logits summed, then K4 == fused K3+K4."""
    B, P, C, H = 1, 500, 6, 4
    iv, cv = rnd(B, H, P, 128, seed=4, scale=0.3).cuda(), rnd(B, H, C, 128, seed=5, scale=0.3).cuda()
    fused = torch.zeros(B, C, P, device="cuda")
    ops.heatmap_accumulate(iv, cv, fused, "softmax", per_head=True)
    total = torch.zeros(B, C, P, device="cuda")
    for h0 in range(0, H, 2):                       # two "ranks" of two heads each
        total += ops.heatmap_logits(iv[:, h0:h0 + 2], cv[:, h0:h0 + 2], per_head=True)
    split = torch.zeros(B, C, P, device="cuda")
    ops.norm_accumulate(total, split, "softmax")
    torch.testing.assert_close(split, fused, rtol=1e-4, atol=1e-5)


def test_round_bf16_emulation(ops):
    """round_bf16 reproduces the reference's bf16 rounding points (logits, probs, mean)."""
    B, P, C, K = 1, 256, 4, 512
    iv, cv = rnd(B, P, K, seed=6), rnd(B, C, K, seed=7, scale=0.1)
    acc = torch.zeros(B, C, P, device="cuda")
    ops.heatmap_accumulate(iv.cuda(), cv.cuda(), acc, "softmax", round_bf16=True)
    heat = ops.finalize_heatmaps(acc, 1, round_bf16=True).cpu().reshape(B, C, P)
    logits = torch.einsum("bpd,bcd->bcp", iv.float(), cv.float()).to(BF16)
    ref = torch.softmax(logits.float(), dim=1).to(BF16).float()
    agree = (heat == ref).float().mean().item()
    assert agree > 0.99, agree                      # fp32 sum order can flip a last bf16 bit


def test_concept_linear_normalize(ops):
    con = rnd(5, 384, seed=8)
    out = ops.concept_linear_normalize(con.cuda()).cpu()
    ref = oracle.linear_normalization(con, dim=-2)           # bf16 arithmetic like the reference
    torch.testing.assert_close(out.float(), ref.float(), rtol=2e-2, atol=2e-3)


# ---------------------------------------------------------------- K5
def test_masks_bit_exact(ops, golden_dir):
    g = np.load(os.path.join(golden_dir, "masks.npz"))
    coeff = gi.mask_inputs()                                  # many exact ties
    am = ops.argmax_mask(coeff[None].cuda().contiguous()).cpu()
    assert am.dtype == torch.int64 and np.array_equal(am[0].numpy(), g["argmax"])
    bm = ops.binary_masks(coeff.cuda().contiguous()).cpu().numpy()
    assert np.array_equal(bm, g["masks_plain"])
    bb = ops.binary_masks(coeff.cuda().contiguous(), apply_blur=True).cpu().numpy()
    assert (bb == g["masks_blur"]).mean() >= 0.999
    blur = ops.gaussian_blur3(coeff.cuda().contiguous()).cpu()
    torch.testing.assert_close(blur, oracle.gaussian_blur3(coeff), rtol=1e-5, atol=1e-6)
    am_bg = ops.argmax_mask(coeff[None].cuda().contiguous(), num_background=2).cpu()[0]
    assert torch.equal(am_bg, oracle.map_background_to_zero(oracle.argmax_mask(coeff), 2))


def test_masks_full_size(ops):
    heat = torch.rand(2, 32, 128, 128, generator=torch.Generator().manual_seed(9))
    heat = torch.round(heat * 64) / 64
    am = ops.argmax_mask(heat.cuda()).cpu()
    assert torch.equal(am, heat.argmax(dim=1))
    bm = ops.binary_masks(heat.cuda()).cpu()
    assert (bm == oracle.binary_masks(heat)).float().mean() >= 0.999


# ---------------------------------------------------------------- host-buffer C ABI
def test_heatmaps_from_vectors_host_matches_reference(golden_dir):
    from conceptattention_b200 import _cabi
    lib = _cabi.load()
    g = np.load(os.path.join(golden_dir, "heatmaps.npz"))
    for tag, (iv, cv) in gi.heatmap_inputs().items():
        per_head = iv.ndim == 6
        ivb, cvb = iv.to(BF16).contiguous(), cv.to(BF16).contiguous()
        Nt, L, B = iv.shape[:3]
        H, P, d = (iv.shape[3], iv.shape[4], iv.shape[5]) if per_head else (1, iv.shape[3], iv.shape[4])
        C = cv.shape[-2]
        heat = torch.empty(B, C, P, dtype=torch.float32)
        ts = (ctypes.c_int * len(gi.HM_TIMESTEPS))(*gi.HM_TIMESTEPS)
        ls = (ctypes.c_int * len(gi.HM_LAYERS))(*gi.HM_LAYERS)
        rc = lib.ca_heatmaps_from_vectors_host(ivb.data_ptr(), cvb.data_ptr(), Nt, L, B, H, P, C, d,
                                               int(per_head), ts, len(ts), ls, len(ls), 0, 0,
                                               heat.data_ptr())
        _cabi.check(rc, "ca_heatmaps_from_vectors_host")
        ref_bf16_inputs = oracle.heatmaps_from_vectors(ivb.float(), cvb.float(), gi.HM_LAYERS,
                                                       gi.HM_TIMESTEPS).reshape(B, C, P)
        torch.testing.assert_close(heat, ref_bf16_inputs, rtol=1e-4, atol=1e-5)
        golden = torch.from_numpy(g[tag + "_softmax"]).reshape(B, C, P)   # reference, fp32 inputs
        assert ((heat - golden).abs() / golden.clamp_min(1e-3)).max() < 5e-2


@pytest.mark.parametrize("B,P,C,H,T,pairs", [(1, 4096, 4, 24, 256, False), (2, 1000, 20, 24, 8, True),
                                             (1, 300, 32, 8, 4, False), (1, 200, 3, 2, 0, True)])
def test_both_spaces_in_one_launch_equal_two_launches(ops, B, P, C, H, T, pairs):
    """ca_heatmap_accumulate_both (output space + cross-attention space of one layer, gridDim.z = 2)
    against the two single-space launches it replaces: accumulators and per-pair slots, the
    per-head view of the packed q_pre, a vector length the TMA kernel does not take (H = 2 -> SIMT)."""
    D = H * 128
    img, con = rnd(B, P, D, seed=21).cuda(), rnd(B, C, D, seed=22, scale=3.0 / math.sqrt(D)).cuda()
    q_pre = rnd(B, H, P + T + C, 128, seed=23, scale=0.3).cuda()
    qi, qc = q_pre[:, :, :P], q_pre[:, :, P + T:]
    acc_a = torch.full((2, B, C, P), 0.25, device="cuda")
    acc_b = acc_a.clone()
    slots = torch.zeros(6, B, C, P, device="cuda") if pairs else None
    slots_b = None if slots is None else slots.clone()
    slot = torch.full((1,), 2, device="cuda", dtype=torch.int32) if pairs else None
    for norm in ("softmax", "sparsemax"):
        ops.heatmap_accumulate(img, con, acc_a[0], norm, pair=slots, pair_slot=slot, pair_slot_off=0)
        ops.heatmap_accumulate(qi, qc, acc_a[1], norm, per_head=True, pair=slots, pair_slot=slot, pair_slot_off=1)
        ops.heatmap_accumulate_both(img, con, acc_b[0], qi, qc, acc_b[1], norm, pair=slots_b, pair_slot=slot,
                                    pair_slot_off=(0, 1))
    # same kernel, but twice the CTAs per launch -> a smaller split-K factor -> another fp32 summation order
    torch.testing.assert_close(acc_b, acc_a, rtol=5e-5, atol=2e-5)
    if pairs:
        torch.testing.assert_close(slots_b, slots, rtol=5e-5, atol=2e-5)
        assert slots_b[2].abs().sum() > 0 and slots_b[3].abs().sum() > 0 and slots_b[0].abs().sum() == 0
