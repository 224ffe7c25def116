"""Tensor-level wrappers over the C ABI.  torch is used for device memory and the
current stream only; every op below is one call into libconceptattention_b200.so.
"""
from __future__ import annotations

import math

import os

import torch

from . import _cabi
from ._cabi import NORM
from ._cabi import check as _check


def check(status, what, n_launch=1):
    global LAUNCHES
    LAUNCHES += n_launch
    _check(status, what)

BF16 = torch.bfloat16

# Instrumentation used by bench.py: number of C-ABI kernel launches, (when ATTN_EVENTS is a
# list) a CUDA-event pair around every K2 launch on its stream, and (when HBM_EVENTS is a dict)
# the same around every K1a / K1b / K1c / K3+K4 launch together with its ALGORITHMIC bytes
# (SURVEY.md section 8d): {kernel: [(event pair, bytes), ...]}.
LAUNCHES = 0
ATTN_EVENTS = None
HBM_EVENTS = None


class _timed:
    """with _timed("K1b", nbytes): launch  — no-op unless bench.py asked for HBM_EVENTS."""

    def __init__(self, name, nbytes):
        self.name, self.nbytes = name, nbytes

    def __enter__(self):
        if HBM_EVENTS is not None:
            self.ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
            self.ev[0].record()

    def __exit__(self, *exc):
        if HBM_EVENTS is not None and exc[0] is None:
            self.ev[1].record()
            HBM_EVENTS.setdefault(self.name, []).append((self.ev, self.nbytes))
        return False


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def _req(t, dtype, name):
    if not (isinstance(t, torch.Tensor) and t.is_cuda):
        raise ValueError(f"{name} must be a CUDA tensor (no CPU path exists)")
    if t.dtype != dtype:
        raise ValueError(f"{name} must be {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise ValueError(f"{name} must be contiguous")
    return t


def _req_rows(t, dtype, name):
    """2-D, unit inner stride; rows may be a column slice of a wider matrix."""
    if not (isinstance(t, torch.Tensor) and t.is_cuda):
        raise ValueError(f"{name} must be a CUDA tensor (no CPU path exists)")
    if t.dtype != dtype or t.ndim != 2 or t.stride(1) != 1:
        raise ValueError(f"{name} must be a 2-D {dtype} tensor with unit inner stride")
    return t


def ln_modulate(x, mod0, mod1=None, split=None, shift_off=0, scale_off=None, out=None):
    """(1+scale)*LN(x)+shift.  x [B,N,D] bf16; mod0/mod1 [B,M] bf16 rows holding the
    modulation chunks; rows >= split use mod1."""
    B, N, D = x.shape
    _req(x, BF16, "x"); _req_rows(mod0, BF16, "mod0")
    if mod1 is not None:
        _req_rows(mod1, BF16, "mod1")
        assert mod1.shape == mod0.shape and mod1.stride(0) == mod0.stride(0)
    if out is None:
        out = torch.empty_like(x)
    scale_off = D if scale_off is None else scale_off
    with _timed(f"K1a ln_modulate rows={B * N}", 4 * B * N * D):
        check(_cabi.load().ca_ln_modulate(
            x.data_ptr(), out.data_ptr(), B, N, D, mod0.data_ptr(),
            mod1.data_ptr() if mod1 is not None else None, N if split is None else split,
            mod0.stride(0), shift_off, scale_off, _stream(x)), "ca_ln_modulate")
    return out


def gate_residual_ln_modulate(x, y, mod0, mod1=None, split=None, gate_off=0, shift_off=0,
                              scale_off=0, out=None, want_out=True):
    """x += gate*y in place; returns (1+scale)*LN(x)+shift if want_out else None.
    y may be a column slice of a wider matrix (row stride y.stride(-2))."""
    B, N, D = x.shape
    _req(x, BF16, "x"); _req_rows(mod0, BF16, "mod0")
    if mod1 is not None:
        _req_rows(mod1, BF16, "mod1")
        assert mod1.stride(0) == mod0.stride(0)
    assert y.dtype == BF16 and y.is_cuda and y.stride(-1) == 1 and y.shape == x.shape
    assert y.stride(0) == N * y.stride(1) or B == 1
    if want_out and out is None:
        out = torch.empty_like(x)
    with _timed(f"K1c gate_residual{'_ln_modulate' if want_out else ''} rows={B * N}", (8 if want_out else 6) * B * N * D):
        check(_cabi.load().ca_gate_residual_ln_modulate(
            x.data_ptr(), y.data_ptr(), y.stride(1), out.data_ptr() if want_out else None, B, N, D,
            mod0.data_ptr(), mod1.data_ptr() if mod1 is not None else None,
            N if split is None else split, mod0.stride(0), gate_off, shift_off, scale_off,
            _stream(x)), "ca_gate_residual_ln_modulate")
    return out if want_out else None


def qkv_prep(qkv_a, qkv_b, scales_a, scales_b, rope, H, want_q_pre=False):
    """Split + QK-RMSNorm + RoPE + head-major layout.
    qkv_a [B,Na,>=3*H*128] (row-strided ok), qkv_b [B,Nb,...] or None;
    scales_* = (q_scale, k_scale) bf16 [128]; rope fp32 [Na+Nb,64,2].
    Returns q,k,v [B,H,Na+Nb,128] (+ q_pre)."""
    B, Na = qkv_a.shape[0], qkv_a.shape[1]
    Nb = 0 if qkv_b is None else qkv_b.shape[1]
    for t in (qkv_a, qkv_b):
        if t is not None:
            assert t.is_cuda and t.dtype == BF16 and t.stride(-1) == 1
            assert t.shape[0] == 1 or t.stride(0) == t.shape[1] * t.stride(1)
    _req(rope, torch.float32, "rope")
    assert rope.shape == (Na + Nb, 64, 2), rope.shape
    dev = qkv_a.device
    q, k, v = (torch.empty(B, H, Na + Nb, 128, device=dev, dtype=BF16) for _ in range(3))
    q_pre = torch.empty_like(q) if want_q_pre else None
    sb = scales_b if scales_b is not None else (None, None)
    Ntok = B * (Na + Nb)
    with _timed(f"K1b qkv_prep{'+q_pre' if want_q_pre else ''} rows={Ntok}", Ntok * (2 * 3 * H * 128 * 2 + 512 + (2 * H * 128 if want_q_pre else 0))):
        check(_cabi.load().ca_qkv_prep(
            qkv_a.data_ptr(), qkv_a.stride(1), Na,
            qkv_b.data_ptr() if qkv_b is not None else None,
            qkv_b.stride(1) if qkv_b is not None else 0, Nb,
            scales_a[0].data_ptr(), scales_a[1].data_ptr(),
            sb[0].data_ptr() if sb[0] is not None else None,
            sb[1].data_ptr() if sb[1] is not None else None,
            rope.data_ptr(), B, H, q.data_ptr(), k.data_ptr(), v.data_ptr(),
            q_pre.data_ptr() if want_q_pre else None, _stream(qkv_a)), "ca_qkv_prep")
    return (q, k, v, q_pre) if want_q_pre else (q, k, v)


def qkv_prep_grouped(qkv_a, qkv_b, scales_a, scales_b, rope, H, G, want_q_pre=False, send=None):
    """K1b for the sequence-parallel exchange (B = 1): same maths as `qkv_prep`, q/k/v written as
    one send buffer [H/G, 3, G, N, 128] (chunk w = the q, k, v of rank w's G heads over this rank's
    N rows).  Returns (send, q_pre | None); q_pre keeps [1, H, N, 128]."""
    B, Na = qkv_a.shape[0], qkv_a.shape[1]
    Nb = 0 if qkv_b is None else qkv_b.shape[1]
    assert B == 1 and H % G == 0
    for t in (qkv_a, qkv_b):
        if t is not None:
            assert t.is_cuda and t.dtype == BF16 and t.stride(-1) == 1
    _req(rope, torch.float32, "rope")
    assert rope.shape == (Na + Nb, 64, 2), rope.shape
    N = Na + Nb
    if send is None:
        send = torch.empty(H // G, 3, G, N, 128, device=qkv_a.device, dtype=BF16)
    else:                                   # caller-owned (e.g. a symmetric-memory region)
        _req(send, BF16, "send")
        assert send.shape == (H // G, 3, G, N, 128) and send.is_contiguous()
    q_pre = torch.empty(1, H, N, 128, device=qkv_a.device, dtype=BF16) if want_q_pre else None
    sb = scales_b if scales_b is not None else (None, None)
    slab = G * N * 128 * 2                                     # bytes of one [G,N,128] slab
    check(_cabi.load().ca_qkv_prep_grouped(
        qkv_a.data_ptr(), qkv_a.stride(1), Na,
        qkv_b.data_ptr() if qkv_b is not None else None,
        qkv_b.stride(1) if qkv_b is not None else 0, Nb,
        scales_a[0].data_ptr(), scales_a[1].data_ptr(),
        sb[0].data_ptr() if sb[0] is not None else None,
        sb[1].data_ptr() if sb[1] is not None else None,
        rope.data_ptr(), 1, H, send.data_ptr(), send.data_ptr() + slab, send.data_ptr() + 2 * slab,
        q_pre.data_ptr() if want_q_pre else None, G, 3 * G * N * 128, _stream(qkv_a)),
        "ca_qkv_prep_grouped")
    return send, q_pre


def seq_unpack_qkv(recv, self_rank, Pl, Tl, C):
    """recv [W,3,Hl,Pl+Tl+C,128] (chunk w from rank w) -> q,k,v [1,Hl,W*(Pl+Tl)+C,128] in K2 row
    order [img | txt | concept]; concept rows come from this rank's own chunk."""
    _req(recv, BF16, "recv")
    W, three, Hl, Nl, d = recv.shape
    assert three == 3 and d == 128 and Nl == Pl + Tl + C
    N = W * (Pl + Tl) + C
    q, k, v = (torch.empty(1, Hl, N, 128, device=recv.device, dtype=BF16) for _ in range(3))
    check(_cabi.load().ca_seq_unpack_qkv(recv.data_ptr(), W, self_rank, Hl, Pl, Tl, C, q.data_ptr(),
                                         k.data_ptr(), v.data_ptr(), _stream(recv)), "ca_seq_unpack_qkv")
    return q, k, v


def seq_gather_qkv(peers_dev, send_off, W, self_rank, Hl, Pl, Tl, C, device):
    """Fused exchange: q,k,v [1,Hl,W*(Pl+Tl)+C,128] pulled out of the peers' send buffers
    (`peers_dev`: device array of W base pointers, `send_off`: byte offset of the send region)."""
    N = W * (Pl + Tl) + C
    q, k, v = (torch.empty(1, Hl, N, 128, device=device, dtype=BF16) for _ in range(3))
    check(_cabi.load().ca_seq_gather_qkv(peers_dev, send_off, W, self_rank, Hl, Pl, Tl, C, q.data_ptr(),
                                         k.data_ptr(), v.data_ptr(),
                                         torch.cuda.current_stream(device).cuda_stream), "ca_seq_gather_qkv")
    return q, k, v


def seq_gather_out(peers_dev, img_off, tc_off, W, self_rank, Hl, Pl, Tl, C, device, out=None):
    """Fused return exchange: this rank's rows with all heads, attn_img [1,Pl,W*Hl*128] and attn_tc
    [1,Tl+C,W*Hl*128], pulled out of every peer's K2 outputs."""
    K = W * Hl * 128
    if out is None:
        attn_img = torch.empty(1, Pl, K, device=device, dtype=BF16)
        attn_tc = torch.empty(1, Tl + C, K, device=device, dtype=BF16)
    else:
        attn_img, attn_tc = out
    check(_cabi.load().ca_seq_gather_out(peers_dev, img_off, tc_off, W, self_rank, Hl, Pl, Tl, C,
                                         attn_img.data_ptr(), attn_tc.data_ptr(),
                                         torch.cuda.current_stream(device).cuda_stream), "ca_seq_gather_out")
    return attn_img, attn_tc


def seq_pack_out(out_img, out_txtc, W, Pl, Tl, C):
    """K2 outputs [1,W*Pl,Hl*128], [1,W*Tl+C,Hl*128] -> send [W,Pl+Tl+C,Hl*128] (every chunk also
    carries the C concept rows of this rank's heads)."""
    _req(out_img, BF16, "out_img")
    _req(out_txtc, BF16, "out_txtc")
    Hl = out_img.shape[-1] // 128
    assert out_img.shape[1] == W * Pl and out_txtc.shape[1] == W * Tl + C
    send = torch.empty(W, Pl + Tl + C, Hl * 128, device=out_img.device, dtype=BF16)
    check(_cabi.load().ca_seq_pack_out(out_img.data_ptr(), out_txtc.data_ptr(), W, Hl, Pl, Tl, C,
                                       send.data_ptr(), _stream(out_img)), "ca_seq_pack_out")
    return send


def seq_unpack_out(recv, Pl, Tl, C, out=None):
    """recv [W,Pl+Tl+C,Hl*128] (chunk w = heads of rank w) -> attn_img [1,Pl,W*Hl*128],
    attn_tc [1,Tl+C,W*Hl*128]: this rank's rows with all heads."""
    _req(recv, BF16, "recv")
    W, Nl, K = recv.shape
    assert Nl == Pl + Tl + C and K % 128 == 0
    if out is not None:                      # caller-owned, dense rows of W*K elements
        attn_img, attn_tc = out
        for t, n in ((attn_img, Pl), (attn_tc, Tl + C)):
            assert t.dtype == BF16 and t.shape[-2:] == (n, W * K) and t.stride(-1) == 1 and t.stride(-2) == W * K
    else:
        attn_img = torch.empty(1, Pl, W * K, device=recv.device, dtype=BF16)
        attn_tc = torch.empty(1, Tl + C, W * K, device=recv.device, dtype=BF16)
    check(_cabi.load().ca_seq_unpack_out(recv.data_ptr(), W, K // 128, Pl, Tl, C,
                                         attn_img.data_ptr() if Pl else None,
                                         attn_tc.data_ptr() if Tl + C else None, _stream(recv)),
          "ca_seq_unpack_out")
    return attn_img, attn_tc


# A/B runs of whole pipelines (bench.py): CA_K2_IMPL=<CA_K2_* id> makes every call that does not name
# a kernel use that one.  Read once at import, passed per call; the library itself has no switches.
_DEFAULT_K2_IMPL = int(os.environ.get("CA_K2_IMPL", str(_cabi.K2_AUTO)))


def joint_attention(q, k, v, P, T, C, mode=_cabi.ATTN_BOTH, out_img=None, out_txtc=None,
                    impl=None, plan_sms=0):
    """One-way joint attention over [img|txt|concept]; returns (out_img [B,P,H*128],
    out_txtc [B,T+C,H*128]).  impl / plan_sms: the per-call kernel and stream-K plan arguments of
    ca_joint_attention_ex (K2_AUTO / 0 = what the product uses)."""
    for n, t in (("q", q), ("k", k), ("v", v)):
        _req(t, BF16, n)
    B, H, N, d = q.shape
    assert d == 128 and N == P + T + C and k.shape == q.shape and v.shape == q.shape
    if impl is None:
        impl = _DEFAULT_K2_IMPL
    if out_img is None:
        out_img = torch.empty(B, P, H * 128, device=q.device, dtype=BF16)
    if out_txtc is None:
        out_txtc = torch.empty(B, T + C, H * 128, device=q.device, dtype=BF16)
    lib = _cabi.load()
    # scratch for the stream-K tail (0 bytes when this shape is never split); torch owns it
    ws_bytes = lib.ca_joint_attention_workspace_bytes_ex(B, H, P, T, C, mode, impl, plan_sms)
    ws = torch.empty(ws_bytes, device=q.device, dtype=torch.uint8) if ws_bytes else None
    ev = None
    if ATTN_EVENTS is not None:
        ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        ev[0].record()
    check(lib.ca_joint_attention_ex(
        q.data_ptr(), k.data_ptr(), v.data_ptr(), B, H, P, T, C, mode,
        out_img.data_ptr() if P > 0 else None, out_txtc.data_ptr() if T + C > 0 else None,
        ws.data_ptr() if ws is not None else None, ws_bytes, impl, plan_sms, _stream(q)), "ca_joint_attention_ex")
    if ev is not None:
        ev[1].record()
        ATTN_EVENTS.append((ev, B, P, T, C, H))
    return out_img, out_txtc


def _hm_strides(img, con, per_head):
    if per_head:      # img [B,H,P,d] view, con [B,H,C,d] view (last dim contiguous)
        B, H, P, d = img.shape
        C = con.shape[2]
        assert img.stride(3) == 1 and con.stride(3) == 1
        return (B, P, C, H, d, (img.stride(0), img.stride(1), img.stride(2)),
                (con.stride(0), con.stride(1), con.stride(2)))
    B, P, K = img.shape
    C = con.shape[1]
    assert img.stride(2) == 1 and con.stride(2) == 1
    return B, P, C, 1, K, (img.stride(0), 0, img.stride(1)), (con.stride(0), 0, con.stride(1))


def heatmap_accumulate(img, con, acc, norm="softmax", round_bf16=False, per_head=False,
                       pair=None, pair_slot=None, pair_slot_off=0):
    """acc[b,c,p] += norm_c(<img[b,p,:], con[b,c,:]>).  Views with arbitrary
    batch/row strides are accepted (no copies).
    `pair` (fp32 [S,B,C,P] slots, optional): the same pass also stores this pair's normalised
    map into slot `pair_slot[0] + pair_slot_off` (`pair_slot`: device int32 tensor or None)."""
    assert img.dtype == BF16 and con.dtype == BF16 and img.is_cuda
    B, P, C, nseg, seg_len, si, sc = _hm_strides(img, con, per_head)
    _req(acc, torch.float32, "acc")
    assert acc.shape == (B, C, P)
    K = nseg * seg_len
    tm = _timed(f"K3+K4 heatmap_accumulate {'cross' if per_head else 'output'} space rows={B * P}", B * (2 * P * K + 2 * C * K + (8 if pair is None else 12) * C * P))
    if pair is None:
        with tm:
            check(_cabi.load().ca_heatmap_accumulate(
                img.data_ptr(), si[0], si[1], si[2], con.data_ptr(), sc[0], sc[1], sc[2], nseg, seg_len,
                B, P, C, NORM[norm], int(round_bf16), acc.data_ptr(), _stream(img)), "ca_heatmap_accumulate")
        return acc
    _req(pair, torch.float32, "pair")
    assert pair.ndim == 4 and pair.shape[1:] == (B, C, P)
    if pair_slot is not None:
        assert pair_slot.is_cuda and pair_slot.dtype == torch.int32 and pair_slot.numel() == 1
    else:
        assert 0 <= pair_slot_off < pair.shape[0]
    with tm:
        check(_cabi.load().ca_heatmap_accumulate_pair(
            img.data_ptr(), si[0], si[1], si[2], con.data_ptr(), sc[0], sc[1], sc[2], nseg, seg_len,
            B, P, C, NORM[norm], int(round_bf16), acc.data_ptr(), pair.data_ptr(), pair.stride(0),
            pair_slot.data_ptr() if pair_slot is not None else None, pair_slot_off, _stream(img)),
            "ca_heatmap_accumulate_pair")
    return acc


def heatmap_logits(img, con, per_head=False, out=None):
    assert img.dtype == BF16 and con.dtype == BF16 and img.is_cuda
    B, P, C, nseg, seg_len, si, sc = _hm_strides(img, con, per_head)
    if out is None:
        out = torch.empty(B, C, P, device=img.device, dtype=torch.float32)
    check(_cabi.load().ca_heatmap_logits(
        img.data_ptr(), si[0], si[1], si[2], con.data_ptr(), sc[0], sc[1], sc[2], nseg, seg_len,
        B, P, C, out.data_ptr(), _stream(img)), "ca_heatmap_logits")
    return out


def heatmap_accumulate_both(img0, con0, acc0, img1, con1, acc1, norm="softmax", round_bf16=False,
                            pair=None, pair_slot=None, pair_slot_off=(0, 1)):
    """Both spaces of one layer in one launch: (img0 [B,P,K], con0 [B,C,K]) -> acc0 and the per-head
    views (img1 [B,H,P,d], con1 [B,H,C,d]) -> acc1; same maths as two `heatmap_accumulate` calls."""
    B, P, C, n0, l0, si0, sc0 = _hm_strides(img0, con0, False)
    B1, P1, C1, n1, l1, si1, sc1 = _hm_strides(img1, con1, True)
    assert (B, P, C) == (B1, P1, C1) and n0 * l0 == n1 * l1
    for a in (acc0, acc1):
        _req(a, torch.float32, "acc")
        assert a.shape == (B, C, P)
    pstride = 0
    if pair is not None:
        _req(pair, torch.float32, "pair")
        assert pair.dim() == 4 and pair.shape[1:] == (B, C, P)
        pstride = pair.stride(0)
    K = n0 * l0
    with _timed(f"K3+K4 heatmap_accumulate both spaces rows={B * P}",
                2 * B * (2 * P * K + 2 * C * K + (8 if pair is None else 12) * C * P)):
        check(_cabi.load().ca_heatmap_accumulate_both(
            img0.data_ptr(), *si0, con0.data_ptr(), *sc0, n0, l0, acc0.data_ptr(), pair_slot_off[0],
            img1.data_ptr(), *si1, con1.data_ptr(), *sc1, n1, l1, acc1.data_ptr(), pair_slot_off[1],
            B, P, C, NORM[norm], int(round_bf16), pair.data_ptr() if pair is not None else None, pstride,
            pair_slot.data_ptr() if pair_slot is not None else None, _stream(img0)),
            "ca_heatmap_accumulate_both")


def norm_accumulate(logits, acc, norm="softmax", round_bf16=False, pair=None):
    """acc += norm_c(logits); `pair` (fp32 [B,C,P], optional) also receives norm_c(logits)."""
    _req(logits, torch.float32, "logits"); _req(acc, torch.float32, "acc")
    B, C, P = logits.shape
    if pair is not None:
        _req(pair, torch.float32, "pair")
        assert pair.shape == (B, C, P)
    check(_cabi.load().ca_norm_accumulate_pair(
        logits.data_ptr(), B, P, C, NORM[norm], int(round_bf16), acc.data_ptr(),
        pair.data_ptr() if pair is not None else None, _stream(logits)), "ca_norm_accumulate_pair")
    return acc


def concept_linear_normalize(con):
    _req(con, BF16, "con")
    C, K = con.shape
    out = torch.empty_like(con)
    check(_cabi.load().ca_concept_linear_normalize(con.data_ptr(), out.data_ptr(), C, K,
                                                   _stream(con)), "ca_concept_linear_normalize")
    return out


def finalize_heatmaps(acc, count, round_bf16=False):
    """mean over the `count` accumulated (t,l) pairs -> [B,C,h,w] fp32."""
    _req(acc, torch.float32, "acc")
    B, C, P = acc.shape
    heat = torch.empty_like(acc)
    check(_cabi.load().ca_finalize_heatmaps(acc.data_ptr(), heat.data_ptr(), acc.numel(),
                                            1.0 / count, int(round_bf16), _stream(acc)),
          "ca_finalize_heatmaps")
    side = math.isqrt(P)
    return heat.view(B, C, side, side) if side * side == P else heat


def gaussian_blur3(maps):
    _req(maps, torch.float32, "maps")
    h, w = maps.shape[-2:]
    out = torch.empty_like(maps)
    check(_cabi.load().ca_gaussian_blur3(maps.data_ptr(), out.data_ptr(), maps.numel() // (h * w),
                                         h, w, _stream(maps)), "ca_gaussian_blur3")
    return out


def argmax_mask(heat, num_background=0, apply_blur=False):
    """heat [B,C,h,w] fp32 -> int64 [B,h,w]; first index on ties."""
    _req(heat, torch.float32, "heat")
    if apply_blur:
        heat = gaussian_blur3(heat)
    B, C, h, w = heat.shape
    mask = torch.empty(B, h, w, device=heat.device, dtype=torch.int64)
    check(_cabi.load().ca_argmax_mask(heat.data_ptr(), mask.data_ptr(), B, C, h * w,
                                      num_background, _stream(heat)), "ca_argmax_mask")
    return mask


def binary_masks(heat, apply_blur=False):
    """heat [...,h,w] fp32 -> bool, coeff > mean_hw(coeff) per map."""
    _req(heat, torch.float32, "heat")
    if apply_blur:
        heat = gaussian_blur3(heat)
    h, w = heat.shape[-2:]
    mask = torch.empty(heat.shape, device=heat.device, dtype=torch.uint8)
    check(_cabi.load().ca_binary_masks(heat.data_ptr(), mask.data_ptr(), heat.numel() // (h * w),
                                       h * w, _stream(heat)), "ca_binary_masks")
    return mask.bool()
