"""GPU parity tests, kernel by kernel: every call goes through the C ABI
(conceptattention_b200.ops -> libconceptattention_b200.so) and is compared with
the CPU oracle on the same seeded inputs.  Run on the B200 box: pytest -m gpu."""
import math

import numpy as np
import pytest
import torch

import oracle

pytestmark = pytest.mark.gpu
BF16 = torch.bfloat16


@pytest.fixture(scope="module")
def ops():
    assert torch.cuda.is_available(), "GPU tests need a B200"
    from conceptattention_b200 import _cabi, ops as _ops
    _cabi.check(_cabi.load().ca_device_check(0), "device")
    return _ops


def rnd(*shape, seed=0, scale=1.0, dtype=BF16):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(dtype)


def close(a, b, rtol, atol):
    torch.testing.assert_close(a.float().cpu(), b.float().cpu(), rtol=rtol, atol=atol)


# ---------------------------------------------------------------- K1a / K1c
@pytest.mark.parametrize("B,N,D,split", [(1, 75, 256, 72), (2, 33, 3072, 30), (1, 260, 3072, 256),
                                         (1, 5, 1024, None)])
def test_ln_modulate(ops, B, N, D, split):
    x = rnd(B, N, D, seed=1, scale=2.0)
    mod0, mod1 = rnd(B, 6 * D, seed=2, scale=0.5), rnd(B, 6 * D, seed=3, scale=0.5)
    out = ops.ln_modulate(x.cuda(), mod0.cuda(), None if split is None else mod1.cuda(), split, 0, D)
    s = N if split is None else split
    xf = x.float()
    ref = torch.empty_like(xf)
    for m, sl in ((mod0, slice(0, s)), (mod1, slice(s, N))):
        mf = m.float()[:, None, :]
        ref[:, sl] = oracle.layer_norm_modulate(xf[:, sl], mf[..., :D], mf[..., D:2 * D])
    close(out, ref, rtol=1e-2, atol=2e-2)


@pytest.mark.parametrize("B,N,D,split,want_out", [(1, 75, 256, 72, True), (2, 40, 3072, 36, True),
                                                  (1, 20, 3072, None, False)])
def test_gate_residual_ln_modulate(ops, B, N, D, split, want_out):
    x, y = rnd(B, N, D, seed=4), rnd(B, N, D, seed=5)
    mod0, mod1 = rnd(B, 6 * D, seed=6, scale=0.5), rnd(B, 6 * D, seed=7, scale=0.5)
    xg = x.cuda().clone()
    out = ops.gate_residual_ln_modulate(xg, y.cuda(), mod0.cuda(),
                                        None if split is None else mod1.cuda(), split,
                                        2 * D, 3 * D, 4 * D, want_out=want_out)
    s = N if split is None else split
    xr = x.float().clone()
    ref = torch.empty_like(xr)
    for m, sl in ((mod0, slice(0, s)), (mod1, slice(s, N))):
        mf = m.float()[:, None, :]
        xr[:, sl] = (xr[:, sl] + mf[..., 2 * D:3 * D] * y.float()[:, sl]).to(BF16).float()
        ref[:, sl] = oracle.layer_norm_modulate(xr[:, sl], mf[..., 3 * D:4 * D], mf[..., 4 * D:5 * D])
    close(xg, xr, rtol=1e-2, atol=1e-2)
    if want_out:
        close(out, ref, rtol=2e-2, atol=3e-2)
    else:
        assert out is None


@pytest.mark.parametrize("B,N,D,split", [(1, 4096, 3072, None), (2, 1111, 3072, 1000), (3, 700, 256, 650),
                                         (1, 1025, 1032, 7)])
def test_ln_streaming_kernels_match_row_kernels_and_oracle(ops, B, N, D, split):
    """Large token streams take the persistent TMA-staged kernels (residual forms, rows >= 1024);
    they must agree BIT FOR BIT with the one-warp-per-row kernels (same summation order) — reached
    here by feeding the same rows in chunks of 256, which the launcher routes to the row kernels —
    and with the oracle: ragged row counts per CTA/warp, B > 1, the per-row modulation switch at
    `split`, a strided y (column slice of a wider matrix), and the no-LN form."""
    x, yw = rnd(B, N, D, seed=11, scale=2.0).cuda(), rnd(B, N, 2 * D, seed=12).cuda()
    y = yw[:, :, D:]
    mod0, mod1 = rnd(B, 6 * D, seed=13, scale=0.5).cuda(), rnd(B, 6 * D, seed=14, scale=0.5).cuda()

    def run(xs, ys, sp):
        a = ops.ln_modulate(xs, mod0, None if sp is None else mod1, sp, 0, D)
        xb = xs.clone()
        b = ops.gate_residual_ln_modulate(xb, ys, mod0, None if sp is None else mod1, sp, 2 * D, 3 * D, 4 * D)
        xc = xs.clone()
        ops.gate_residual_ln_modulate(xc, ys, mod0, None if sp is None else mod1, sp, 5 * D, 0, 0, want_out=False)
        return a, b, xb, xc

    whole = run(x, y, split)
    parts = []
    for a0 in range(0, N, 256):
        b0 = min(N, a0 + 256)
        sp = None if split is None else min(max(split - a0, 0), b0 - a0)
        parts.append(run(x[:, a0:b0].contiguous(), yw[:, a0:b0].contiguous()[:, :, D:], sp))
    torch.cuda.synchronize()
    for i, got in enumerate(whole):
        assert torch.equal(got, torch.cat([pt[i] for pt in parts], dim=1)), i
    s_ = N if split is None else split
    xf = x.float().cpu()
    ref = torch.empty_like(xf)
    for m, sl in ((mod0, slice(0, s_)), (mod1, slice(s_, N))):
        mf = m.float().cpu()[:, None, :]
        ref[:, sl] = oracle.layer_norm_modulate(xf[:, sl], mf[..., :D], mf[..., D:2 * D])
    close(whole[0], ref, rtol=1e-2, atol=2e-2)


# ---------------------------------------------------------------- K1b
@pytest.mark.parametrize("B,Na,Nb,H,extra", [(1, 144, 75, 2, 0), (2, 64, 20, 3, 0), (1, 100, 0, 2, 512)])
def test_qkv_prep(ops, B, Na, Nb, H, extra):
    D = H * 128
    qa = rnd(B, Na, 3 * D + extra, seed=8)
    qb = rnd(B, Nb, 3 * D, seed=9) if Nb else None
    sc = [(1 + 0.3 * torch.randn(128, generator=torch.Generator().manual_seed(10 + i))).to(BF16)
          for i in range(4)]
    side = int(math.ceil(math.sqrt(Na)))
    ids = torch.zeros(Na + Nb, 3)
    ids[:Na, 1] = torch.arange(Na) // side
    ids[:Na, 2] = torch.arange(Na) % side
    if Nb:
        ids[Na:Na + 3, 1] = 5.0        # non-zero text ids must rotate too (general path)
    pe = oracle.rope_table(ids[None], [16, 56, 56], 10000)                  # [1,1,N,64,2,2]
    table = torch.stack([pe[0, 0, :, :, 0, 0], pe[0, 0, :, :, 1, 0]], -1).contiguous()   # cos, sin
    res = ops.qkv_prep(qa.cuda(), qb.cuda() if Nb else None, (sc[0].cuda(), sc[1].cuda()),
                       (sc[2].cuda(), sc[3].cuda()) if Nb else None, table.cuda(), H, want_q_pre=True)
    q, k, v, q_pre = [t.cpu() for t in res]
    # oracle, stream by stream
    qs, ks, vs, qps = [], [], [], []
    for src, (s_q, s_k) in ((qa[..., :3 * D], (sc[0], sc[1])),) + (((qb, (sc[2], sc[3])),) if Nb else ()):
        oq, ok, ov = oracle.split_heads(src, H)
        oq, ok = oracle.rms_norm_qk(oq, ok, ov, s_q, s_k)
        qs.append(oq); ks.append(ok); vs.append(ov); qps.append(oq)
    oq, ok, ov, oqp = (torch.cat(t, 2) for t in (qs, ks, vs, qps))
    oq, ok = oracle.apply_rope(oq, ok, pe)
    assert torch.equal(v, ov)
    close(q_pre, oqp, rtol=1e-2, atol=1e-2)
    close(q, oq, rtol=1e-2, atol=2e-2)
    close(k, ok, rtol=1e-2, atol=2e-2)


# ---------------------------------------------------------------- K2
def _attention_oracle(q, k, v, P, T, C, mode):
    """q,k,v [B,H,P+T+C,128] fp32 in packed order -> (img [B,P,HD], txtc [B,T+C,HD])."""
    PT = P + T
    main = oracle.sdpa_explicit(q[:, :, :PT], k[:, :, :PT], v[:, :, :PT])
    outs = [main]
    if C:
        cq = q[:, :, PT:]
        if mode == 0:
            kk, vv = torch.cat((k[:, :, :P], k[:, :, PT:]), 2), torch.cat((v[:, :, :P], v[:, :, PT:]), 2)
            outs.append(oracle.sdpa_explicit(cq, kk, vv))
        elif mode == 1:
            outs.append(oracle.sdpa_explicit(cq, k[:, :, :P], v[:, :, :P]))
        elif mode == 2:
            outs.append(oracle.sdpa_explicit(cq, k[:, :, PT:], v[:, :, PT:]))
        else:
            outs.append(v[:, :, PT:])
    o = oracle.merge_heads(torch.cat(outs, 2))
    return o[:, :P], o[:, P:]


ATTN_CASES = [
    (1, 2, 144, 72, 3, 0), (1, 2, 144, 72, 3, 1), (1, 2, 144, 72, 3, 2), (1, 2, 144, 72, 3, 3),
    (2, 3, 256, 256, 4, 0), (1, 2, 1024, 256, 32, 0), (1, 1, 200, 100, 1, 0), (1, 2, 384, 128, 0, 0),
    (1, 1, 64, 0, 2, 0),
]


@pytest.mark.parametrize("B,H,P,T,C,mode", ATTN_CASES)
def test_joint_attention(ops, B, H, P, T, C, mode):
    N = P + T + C
    q, k, v = rnd(B, H, N, 128, seed=20, scale=1.5), rnd(B, H, N, 128, seed=21), rnd(B, H, N, 128, seed=22)
    oi, ot = ops.joint_attention(q.cuda(), k.cuda(), v.cuda(), P, T, C, mode)
    ri, rt = _attention_oracle(q.float(), k.float(), v.float(), P, T, C, mode)
    close(oi, ri, rtol=2e-2, atol=1e-2)
    close(ot, rt, rtol=2e-2, atol=1e-2)


@pytest.mark.parametrize("B,H,P,T,C,sms", [(1, 2, 2000, 200, 3, 7), (1, 2, 2200, 200, 3, 6), (1, 2, 2200, 200, 0, 6),
                                           (1, 2, 2200, 200, 5, 24), (2, 3, 1100, 100, 2, 8)])
def test_joint_attention_stream_k_tail(ops, B, H, P, T, C, sms):
    """Stream-K tail: with the plan made for `sms` SMs the left-over query-tile units are cut
    by KEY range into one share per SM and merged by the last piece to finish.  The result
    must match the oracle like the unsplit launch does, be (near-)identical to the unsplit
    launch (fp32 merge vs in-TMEM accumulation: one bf16 ulp), and be REPRODUCIBLE — the
    merge order is fixed, so repeated launches are bit-identical whatever finishes last."""
    from conceptattention_b200 import _cabi
    lib = _cabi.load()
    N = P + T + C
    q, k, v = (rnd(B, H, N, 128, seed=s_, scale=sc).cuda() for s_, sc in ((30, 1.5), (31, 1.0), (32, 1.0)))
    V3 = _cabi.K2_V3                                   # the kernel with the stream-K tail, chosen per call
    assert lib.ca_joint_attention_workspace_bytes_ex(B, H, P, T, C, 0, V3, -1) == 0
    base = ops.joint_attention(q, k, v, P, T, C, impl=V3, plan_sms=-1)
    assert lib.ca_joint_attention_workspace_bytes_ex(B, H, P, T, C, 0, V3, sms) > 0, "plan did not split: pick other sizes"
    runs = [ops.joint_attention(q, k, v, P, T, C, impl=V3, plan_sms=sms) for _ in range(3)]
    torch.cuda.synchronize()
    ri, rt = _attention_oracle(q.float().cpu(), k.float().cpu(), v.float().cpu(), P, T, C, 0)
    for got, ref, b in zip(runs[0], (ri, rt), base):
        close(got, ref, rtol=2e-2, atol=1e-2)
        close(got, b, rtol=8e-3, atol=2e-3)
    for r in runs[1:]:
        assert torch.equal(r[0], runs[0][0]) and torch.equal(r[1], runs[0][1])


@pytest.mark.parametrize("impl", ["V3", "V7", "V11", "V11P"])
@pytest.mark.parametrize("B,H,P,T,C,mode", [(1, 2, 1024, 256, 32, 0), (2, 3, 256, 256, 4, 0), (1, 2, 2200, 200, 5, 0),
                                            (1, 1, 200, 100, 1, 0), (1, 2, 144, 72, 3, 1), (1, 2, 100, 20, 3, 2),
                                            (2, 2, 0, 130, 0, 0), (8, 24, 256, 128, 2, 0)])
def test_joint_attention_every_kernel_matches_oracle(ops, impl, B, H, P, T, C, mode):
    """ca_joint_attention_ex with each kernel named per call (CA_K2_V3 / V7 / V11 / V11P): same function to
    bf16 round-off — also with scores that keep growing along the keys (forces the lazy O rescale
    on many tiles), one-tile units, ragged key tiles in both key segments, and more concept
    units than SMs (8 x 24 of them: CTAs that follow a one-tile CTA on the same SM)."""
    from conceptattention_b200 import _cabi
    which = getattr(_cabi, "K2_" + impl)
    N = P + T + C
    q, k, v = (rnd(B, H, N, 128, seed=s_, scale=sc).cuda() for s_, sc in ((40, 1.5), (41, 1.0), (42, 1.0)))
    ramp = torch.linspace(0.2, 6.0, N, device="cuda").view(1, 1, N, 1)          # later keys score higher
    for kk in ((k,) if B * H > 8 else (k, (k.float() * ramp).to(BF16))):
        oi, ot = ops.joint_attention(q, kk, v, P, T, C, mode, impl=which)
        torch.cuda.synchronize()
        ri, rt = _attention_oracle(q.float().cpu(), kk.float().cpu(), v.float().cpu(), P, T, C, mode)
        if P:
            close(oi, ri, rtol=2e-2, atol=1e-2)
        close(ot, rt, rtol=2e-2, atol=1e-2)


@pytest.mark.parametrize("B,H,P,T,C,mode", [(1, 8, 2048, 128, 3, 0), (1, 6, 4096, 512, 20, 0), (2, 10, 1300, 200, 5, 1),
                                            (1, 3, 16384, 256, 32, 0), (1, 24, 4096, 256, 4, 0)])
def test_joint_attention_persistent_key_step_ranges(ops, B, H, P, T, C, mode):
    """CA_K2_V11P with its balanced tail (plan_sms = 1: whenever structurally possible; the default
    plans it when the measured rule says it pays): whole waves of units in lockstep, then every CTA
    works on an equal share of the key steps of the left-over units — all units when there is less
    than one wave, 136 of 408 at the last shape — and units cut by a share boundary are merged in
    fp32 by the piece that finishes last.  Must match the one-unit-per-CTA kernel to a bf16 ulp (fp32 merge vs in-TMEM accumulation),
    the oracle like every kernel, and be REPRODUCIBLE bit for bit (fixed merge order); plan_sms < 0
    turns the ranges off (whole units, no workspace)."""
    from conceptattention_b200 import _cabi
    lib = _cabi.load()
    N = P + T + C
    q, k, v = (rnd(B, H, N, 128, seed=s_, scale=sc).cuda() for s_, sc in ((60, 1.5), (61, 1.0), (62, 1.0)))
    V11P, V11 = _cabi.K2_V11P, _cabi.K2_V11
    assert lib.ca_joint_attention_workspace_bytes_ex(B, H, P, T, C, mode, V11P, 1) > 0, "ranges not planned: pick other sizes"
    assert lib.ca_joint_attention_workspace_bytes_ex(B, H, P, T, C, mode, V11P, -1) == 0
    base = ops.joint_attention(q, k, v, P, T, C, mode, impl=V11)
    whole = ops.joint_attention(q, k, v, P, T, C, mode, impl=V11P, plan_sms=-1)
    runs = [ops.joint_attention(q, k, v, P, T, C, mode, impl=V11P, plan_sms=1) for _ in range(3)]
    auto = ops.joint_attention(q, k, v, P, T, C, mode)
    torch.cuda.synchronize()
    for got, b, w in zip(runs[0], base, whole):
        assert torch.equal(b, w)                            # same per-unit maths without the ranges
        close(got, b, rtol=8e-3, atol=2e-3)
    for r in runs[1:]:
        assert torch.equal(r[0], runs[0][0]) and torch.equal(r[1], runs[0][1])
    if P <= 4096 and H <= 8:
        ri, rt = _attention_oracle(q.float().cpu(), k.float().cpu(), v.float().cpu(), P, T, C, mode)
        close(runs[0][0], ri, rtol=2e-2, atol=1e-2)
        close(runs[0][1], rt, rtol=2e-2, atol=1e-2)
    assert auto[0].shape == runs[0][0].shape


def test_joint_attention_one_way(ops):
    """Image and text rows must not depend on the concept tokens at all (bit-exact)."""
    B, H, P, T, C = 1, 2, 384, 128, 5
    N = P + T + C
    q, k, v = (rnd(B, H, N, 128, seed=s) for s in (30, 31, 32))
    a_i, a_t = ops.joint_attention(q.cuda(), k.cuda(), v.cuda(), P, T, C)
    q2, k2, v2 = q.clone(), k.clone(), v.clone()
    for t in (q2, k2, v2):
        t[:, :, P + T:] = rnd(B, H, C, 128, seed=33, scale=3.0)
    b_i, b_t = ops.joint_attention(q2.cuda(), k2.cuda(), v2.cuda(), P, T, C)
    assert torch.equal(a_i, b_i) and torch.equal(a_t[:, :T], b_t[:, :T])
    assert not torch.equal(a_t[:, T:], b_t[:, T:])


@pytest.mark.parametrize("P,T,C,H", [(4096, 256, 4, 24)])
def test_joint_attention_full_size_properties(ops, P, T, C, H):
    """BASELINE config-2 shape: V = 1 -> every output is 1 (softmax rows sum to 1), and two
    heads are checked against the oracle."""
    N = P + T + C
    q, k = rnd(1, H, N, 128, seed=40), rnd(1, H, N, 128, seed=41)
    ones = torch.ones(1, H, N, 128, dtype=BF16)
    oi, ot = ops.joint_attention(q.cuda(), k.cuda(), ones.cuda(), P, T, C)
    assert (oi.float() - 1).abs().max() < 8e-3 and (ot.float() - 1).abs().max() < 8e-3
    v = rnd(1, H, N, 128, seed=42)
    oi, ot = ops.joint_attention(q.cuda(), k.cuda(), v.cuda(), P, T, C)
    hs = [0, H - 1]
    ri, rt = _attention_oracle(q[:, hs].float(), k[:, hs].float(), v[:, hs].float(), P, T, C, 0)
    for j, h in enumerate(hs):
        close(oi[..., h * 128:(h + 1) * 128], ri[..., j * 128:(j + 1) * 128], rtol=2e-2, atol=5e-3)
        close(ot[..., h * 128:(h + 1) * 128], rt[..., j * 128:(j + 1) * 128], rtol=2e-2, atol=5e-3)


# ---------------------------------------------------------------- C-ABI error behaviour / edges
def test_cabi_status_codes_and_empty_inputs(ops):
    """Error convention of include/conceptattention_b200.h: int status, never throws; empty
    problems are no-ops; shapes outside the kernels' range are refused, not mis-computed."""
    from conceptattention_b200 import _cabi
    lib = _cabi.load()
    q = torch.zeros(1, 2, 40, 128, device="cuda", dtype=BF16)
    oi = torch.zeros(1, 32, 256, device="cuda", dtype=BF16)
    ot = torch.zeros(1, 8, 256, device="cuda", dtype=BF16)
    ptr = lambda t: t.data_ptr()
    call = lambda *a: lib.ca_joint_attention(*a, None)
    assert call(ptr(q), ptr(q), ptr(q), 1, 2, 32, 4, 4, 0, ptr(oi), ptr(ot)) == 0
    assert call(None, ptr(q), ptr(q), 1, 2, 32, 4, 4, 0, ptr(oi), ptr(ot)) == -1          # null pointer
    assert call(ptr(q), ptr(q), ptr(q), 1, 2, 32, 4, 4, 7, ptr(oi), ptr(ot)) == -1        # bad mode
    assert call(ptr(q), ptr(q), ptr(q), 1, 2, 32, 4, 33, 0, ptr(oi), ptr(ot)) == -2       # C > 32
    assert call(ptr(q) + 2, ptr(q), ptr(q), 1, 2, 32, 4, 4, 0, ptr(oi), ptr(ot)) == -3    # misaligned
    assert call(ptr(q), ptr(q), ptr(q), 0, 2, 32, 4, 4, 0, ptr(oi), ptr(ot)) == 0         # empty batch
    assert lib.ca_device_check(0) == 0
    acc = torch.zeros(1, 4, 32, device="cuda")
    st = lambda *a: lib.ca_heatmap_accumulate(*a, None)
    assert st(ptr(oi), 32 * 256, 0, 256, ptr(ot), 8 * 256, 0, 256, 1, 256, 1, 32, 4, 0, 0, ptr(acc)) == 0
    assert st(ptr(oi), 32 * 256, 0, 256, ptr(ot), 8 * 256, 0, 256, 1, 256, 1, 32, 4, 9, 0, ptr(acc)) == -1   # bad norm
    assert st(ptr(oi), 32 * 256, 0, 256, ptr(ot), 8 * 256, 0, 256, 1, 256, 1, 32, 40, 0, 0, ptr(acc)) == -2  # C > 32
    assert st(ptr(oi), 32 * 256, 0, 250, ptr(ot), 8 * 256, 0, 256, 1, 256, 1, 32, 4, 0, 0, ptr(acc)) == -3   # stride % 8
    assert st(ptr(oi), 32 * 256, 0, 256, ptr(ot), 8 * 256, 0, 256, 1, 256, 1, 0, 4, 0, 0, ptr(acc)) == 0     # P = 0
    assert lib.ca_argmax_mask(None, None, 1, 2, 4, 0, None) == -1
    assert lib.ca_status_string(-2).decode() == "unsupported shape"
    torch.cuda.synchronize()
    assert float(acc.abs().sum()) > 0      # the valid call did accumulate (softmax rows sum to 1)
    torch.testing.assert_close(acc.sum(1).cpu(), torch.ones(1, 32))


def test_joint_attention_max_concepts_and_single_concept(ops):
    """C = 32 (CA_MAX_CONCEPTS) and C = 1 through the concept tile; ragged P and T."""
    for C in (1, 32):
        B, H, P, T = 1, 2, 200, 77
        q, k, v = (rnd(B, H, P + T + C, 128, seed=40 + i, scale=0.6) for i in range(3))
        oi, ot = ops.joint_attention(q.cuda(), k.cuda(), v.cuda(), P, T, C)
        qf, kf, vf = q.float(), k.float(), v.float()
        main = oracle.sdpa(qf[:, :, :P + T], kf[:, :, :P + T], vf[:, :, :P + T])
        idx = list(range(P)) + list(range(P + T, P + T + C))
        con = oracle.sdpa(qf[:, :, P + T:], kf[:, :, idx], vf[:, :, idx])
        close(oi, oracle.merge_heads(main[:, :, :P]), rtol=2e-2, atol=2e-2)
        close(ot[:, :T], oracle.merge_heads(main[:, :, P:]), rtol=2e-2, atol=2e-2)
        close(ot[:, T:], oracle.merge_heads(con), rtol=2e-2, atol=2e-2)
