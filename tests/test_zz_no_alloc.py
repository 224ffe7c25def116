"""ca_joint_attention_noalloc (no counterpart in the reference):
K2 with the scratch of its tail balancing in a static device array instead of a
caller-owned workspace, called through the C ABI.  Against the existing entry point
(ca_joint_attention_ex with a workspace torch allocated) on calls in a row that share the array —
the same shape twice and another shape in between, so that stale partials and counters of the last
call are what the next one finds — plus the size limit.  Collected last on purpose (file name)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
BF16 = torch.bfloat16
CA_ERR_SCRATCH_LIMIT = -6
MAX_BYTES = 40 << 20


def _rnd(*shape, seed, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(BF16).cuda()


def _noalloc(lib, q, k, v, P, T, Cn, mode, impl, plan_sms):
    B, H, N, _ = q.shape
    out_img = torch.full((B, P, H * 128), float("nan"), device="cuda", dtype=BF16)
    out_tc = torch.full((B, T + Cn, H * 128), float("nan"), device="cuda", dtype=BF16)
    st = lib.ca_joint_attention_noalloc(
        q.data_ptr(), k.data_ptr(), v.data_ptr(), B, H, P, T, Cn, mode,
        out_img.data_ptr() if P else None, out_tc.data_ptr() if T + Cn else None, impl, plan_sms,
        C.c_void_p(torch.cuda.current_stream().cuda_stream))
    return st, out_img, out_tc


# (B, H, P, T, C, mode, impl, plan_sms): the persistent kernel with its key-step ranges (CA_K2_V11P = 4,
# plan_sms = 1: planned whenever structurally possible),
# round 1's stream-K tail planned for 24 SMs (CA_K2_V3 = 1), and the default on a shape without scratch
CASES = [(1, 8, 2048, 128, 3, 0, 4, 1), (1, 2, 2200, 200, 5, 0, 1, 24), (1, 6, 4096, 512, 20, 0, 4, 1),
         (2, 3, 256, 256, 4, 0, 0, 0)]


def test_noalloc_entry_matches_the_workspace_entry_on_calls_in_a_row():
    from conceptattention_b200 import _cabi, ops
    lib = _cabi.load()
    data, want = [], []
    for i, (B, H, P, T, Cn, mode, impl, sms) in enumerate(CASES):
        N = P + T + Cn
        q, k, v = _rnd(B, H, N, 128, seed=70 + 3 * i, scale=1.5), _rnd(B, H, N, 128, seed=71 + 3 * i), \
            _rnd(B, H, N, 128, seed=72 + 3 * i)
        data.append((q, k, v))
        want.append(ops.joint_attention(q, k, v, P, T, Cn, mode, impl=impl, plan_sms=sms))   # existing entry point
    need = [lib.ca_joint_attention_workspace_bytes_ex(B, H, P, T, Cn, mode, impl, sms)
            for (B, H, P, T, Cn, mode, impl, sms) in CASES]
    assert need[0] > 0 and need[1] > 0 and need[2] > 0 and need[3] == 0 and max(need) <= MAX_BYTES, need
    torch.cuda.synchronize()
    for i in (0, 0, 1, 0, 2, 1, 3, 2):                      # calls in a row on one stream, sharing the array
        B, H, P, T, Cn, mode, impl, sms = CASES[i]
        st, oi, ot = _noalloc(lib, *data[i], P, T, Cn, mode, impl, sms)
        assert st == 0, lib.ca_status_string(st)
        torch.cuda.synchronize()
        for got, ref in ((oi, want[i][0]), (ot, want[i][1])):
            assert not torch.isnan(got.float()).any()
            # the tolerance of the kernel tests; the merge order is fixed, so in fact bit-identical
            torch.testing.assert_close(got.float(), ref.float(), rtol=8e-3, atol=2e-3)
            assert torch.equal(got, ref)


def test_noalloc_entry_rejects_a_plan_beyond_the_static_array():
    from conceptattention_b200 import _cabi
    lib = _cabi.load()
    B, H, P, T, Cn = 1, 24, 4096, 256, 4
    need = lib.ca_joint_attention_workspace_bytes_ex(B, H, P, T, Cn, 0, _cabi.K2_V3, 500)   # planned for 500 SMs
    assert need > MAX_BYTES, need
    q = _rnd(B, H, P + T + Cn, 128, seed=90)
    st, oi, ot = _noalloc(lib, q, q, q, P, T, Cn, 0, _cabi.K2_V3, 500)
    torch.cuda.synchronize()
    assert st == CA_ERR_SCRATCH_LIMIT
    assert b"CA_K2_NOALLOC_MAX_BYTES" in lib.ca_status_string(st)
    assert torch.isnan(oi.float()).all() and torch.isnan(ot.float()).all()      # nothing was launched
    st, oi, ot = _noalloc(lib, q, q, q, P, T, Cn, 0, _cabi.K2_V3, -1)           # never split: no scratch needed
    torch.cuda.synchronize()
    assert st == 0 and not torch.isnan(oi.float()).any()
