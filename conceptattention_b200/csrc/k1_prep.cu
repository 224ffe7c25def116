// K1 family: the memory-bound elementwise/normalisation steps around the cuBLAS
// linears of a double-stream block.  All HBM-bound: one pass, 16-byte accesses,
// fp32 maths, bf16 storage, warp-shuffle reductions.
//
//   ca_ln_modulate                 modified_double_stream_block.py:88-89,94-95,100-101
//   ca_gate_residual_ln_modulate   modified_double_stream_block.py:194-202
//   ca_qkv_prep                    :91-92,97-98,103-104,110,125 + layers.py:63-84 + math.py:25-30
#include "ca_common.cuh"

namespace ca {

constexpr int kWarpsPerCta = 4;

// Row-wise LN + modulate.  One warp per row, the row lives in registers
// (NCH 16-byte chunks per lane), so x is read exactly once.
template <int NCH, bool kResidual>
__global__ void __launch_bounds__(kWarpsPerCta * 32)
ln_modulate_kernel(__nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ y,
                   int64_t y_stride, __nv_bfloat16* __restrict__ out, int rows, int N, int D,
                   const __nv_bfloat16* __restrict__ mod0, const __nv_bfloat16* __restrict__ mod1,
                   int split, int64_t mod_stride, int gate_off, int shift_off, int scale_off) {
  const int row = blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const int b = row / N, n = row - b * N;
  const __nv_bfloat16* mod = (n < split ? mod0 : mod1) + (int64_t)b * mod_stride;
  const int nchunks = D >> 3;
  __nv_bfloat16* xr = x + (int64_t)row * D;

  float v[NCH][8];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < NCH; ++i) {
    const int c = lane + i * 32;
    if (c < nchunks) {
      uint4 u = kResidual ? *reinterpret_cast<const uint4*>(xr + c * 8) : ld_stream16(xr + c * 8);
      unpack8(u, v[i]);
      if (kResidual) {
        float yv[8], gv[8];
        unpack8(ld_stream16(y + (int64_t)row * y_stride + c * 8), yv);
        unpack8(*reinterpret_cast<const uint4*>(mod + gate_off + c * 8), gv);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[i][j] = bf16_round(v[i][j] + gv[j] * yv[j]);
        *reinterpret_cast<uint4*>(xr + c * 8) = pack8(v[i]);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) sum += v[i][j];
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[i][j] = 0.f;
    }
  }
  if (out == nullptr) return;
  const float mean = warp_sum(sum) / D;
  float sq = 0.f;
#pragma unroll
  for (int i = 0; i < NCH; ++i) {
    if (lane + i * 32 < nchunks) {
#pragma unroll
      for (int j = 0; j < 8; ++j) { float d = v[i][j] - mean; sq += d * d; }
    }
  }
  const float rstd = rsqrtf(warp_sum(sq) / D + 1e-6f);
#pragma unroll
  for (int i = 0; i < NCH; ++i) {
    const int c = lane + i * 32;
    if (c < nchunks) {
      float sh[8], sc[8], o[8];
      unpack8(*reinterpret_cast<const uint4*>(mod + shift_off + c * 8), sh);
      unpack8(*reinterpret_cast<const uint4*>(mod + scale_off + c * 8), sc);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = (1.f + sc[j]) * ((v[i][j] - mean) * rstd) + sh[j];
      *reinterpret_cast<uint4*>(out + (int64_t)row * D + c * 8) = pack8(o);
    }
  }
}

template <bool kResidual>
static ca_status launch_ln(void* x, const void* y, int64_t y_stride, void* out, int B, int N, int D,
                           const void* mod0, const void* mod1, int split, int64_t mod_stride,
                           int gate_off, int shift_off, int scale_off, cudaStream_t st) {
  if (!x || !mod0 || B < 0 || N < 0 || D <= 0) return CA_ERR_INVALID_ARG;
  if (!mod1) { mod1 = mod0; }
  if (D % 8 || D > 32 * 8 * 32) return CA_ERR_UNSUPPORTED;
  if (!aligned16(x) || (out && !aligned16(out)) || (y && !aligned16(y)) || !aligned16(mod0) ||
      !aligned16(mod1) || (mod_stride % 8) || (gate_off % 8) || (shift_off % 8) ||
      (scale_off % 8) || (y_stride % 8))
    return CA_ERR_ALIGNMENT;
  const int rows = B * N;
  if (rows == 0) return CA_OK;
  const int nch = (D / 8 + 31) / 32;
  dim3 grid((rows + kWarpsPerCta - 1) / kWarpsPerCta), block(kWarpsPerCta * 32);
#define CA_LN_CASE(NCH)                                                                         \
  ln_modulate_kernel<NCH, kResidual><<<grid, block, 0, st>>>(                                   \
      (__nv_bfloat16*)x, (const __nv_bfloat16*)y, y_stride, (__nv_bfloat16*)out, rows, N, D,    \
      (const __nv_bfloat16*)mod0, (const __nv_bfloat16*)mod1, split, mod_stride, gate_off,      \
      shift_off, scale_off)
  if (nch <= 1) CA_LN_CASE(1);
  else if (nch <= 2) CA_LN_CASE(2);
  else if (nch <= 4) CA_LN_CASE(4);
  else if (nch <= 8) CA_LN_CASE(8);
  else if (nch <= 12) CA_LN_CASE(12);
  else if (nch <= 16) CA_LN_CASE(16);
  else CA_LN_CASE(32);
#undef CA_LN_CASE
  CA_LAUNCH_CHECK();
  return CA_OK;
}

// ---------------------------------------------------------------------------
// qkv split + QK RMSNorm + RoPE + layout.  One warp = one token x two heads;
// a half-warp owns one head (16 lanes x 8 channels).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
qkv_prep_kernel(const __nv_bfloat16* __restrict__ qkv_a, int64_t stride_a, int Na,
                const __nv_bfloat16* __restrict__ qkv_b, int64_t stride_b, int Nb,
                const __nv_bfloat16* __restrict__ qs_a, const __nv_bfloat16* __restrict__ ks_a,
                const __nv_bfloat16* __restrict__ qs_b, const __nv_bfloat16* __restrict__ ks_b,
                const float* __restrict__ rope, int B, int H,
                __nv_bfloat16* __restrict__ q, __nv_bfloat16* __restrict__ k,
                __nv_bfloat16* __restrict__ v, __nv_bfloat16* __restrict__ q_pre) {
  const int hp_per_tok = (H + 1) >> 1;
  const int64_t warp = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int Ntot = Na + Nb;
  if (warp >= (int64_t)B * Ntot * hp_per_tok) return;
  const int lane = threadIdx.x & 31;
  const int hp = (int)(warp % hp_per_tok);
  const int64_t tokg = warp / hp_per_tok;
  const int n = (int)(tokg % Ntot), b = (int)(tokg / Ntot);
  const int h = hp * 2 + (lane >> 4);
  const int ch = lane & 15;                  // 8-channel chunk inside the head
  const bool active = h < H;
  const int hc = active ? h : H - 1;

  const bool in_a = n < Na;
  const __nv_bfloat16* src = in_a ? qkv_a + ((int64_t)b * Na + n) * stride_a
                                  : qkv_b + ((int64_t)b * Nb + (n - Na)) * stride_b;
  const __nv_bfloat16* qs = in_a ? qs_a : qs_b;
  const __nv_bfloat16* ks = in_a ? ks_a : ks_b;
  const int D = H * CA_HEAD_DIM;
  const int col = hc * CA_HEAD_DIM + ch * 8;

  float fq[8], fk[8];
  unpack8(ld_stream16(src + col), fq);
  unpack8(ld_stream16(src + D + col), fk);
  const uint4 uv = ld_stream16(src + 2 * D + col);

  float sq = 0.f, sk = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) { sq += fq[j] * fq[j]; sk += fk[j] * fk[j]; }
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) {          // reduce inside the half-warp
    sq += __shfl_xor_sync(0xffffffffu, sq, o);
    sk += __shfl_xor_sync(0xffffffffu, sk, o);
  }
  const float rq = rsqrtf(sq * (1.f / CA_HEAD_DIM) + 1e-6f);
  const float rk = rsqrtf(sk * (1.f / CA_HEAD_DIM) + 1e-6f);
  float sqv[8], skv[8];
  unpack8(*reinterpret_cast<const uint4*>(qs + ch * 8), sqv);
  unpack8(*reinterpret_cast<const uint4*>(ks + ch * 8), skv);
#pragma unroll
  for (int j = 0; j < 8; ++j) {              // (x*rrms).to(bf16) * scale -> bf16   (layers.py:72)
    fq[j] = bf16_round(bf16_round(fq[j] * rq) * sqv[j]);
    fk[j] = bf16_round(bf16_round(fk[j] * rk) * skv[j]);
  }
  if (!active) return;
  const int64_t dst = (((int64_t)b * H + h) * Ntot + n) * CA_HEAD_DIM + ch * 8;
  if (q_pre) *reinterpret_cast<uint4*>(q_pre + dst) = pack8(fq);
  // RoPE: pairs (2i,2i+1); out0 = cos*x0 - sin*x1, out1 = sin*x0 + cos*x1   (math.py:19-30)
  const float4* rp = reinterpret_cast<const float4*>(rope + ((int64_t)n * 64 + ch * 4) * 2);
  const float4 r0 = rp[0], r1 = rp[1];
  const float cs[4] = {r0.x, r0.z, r1.x, r1.z}, sn[4] = {r0.y, r0.w, r1.y, r1.w};
  float oq[8], ok[8];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    oq[2 * i] = cs[i] * fq[2 * i] - sn[i] * fq[2 * i + 1];
    oq[2 * i + 1] = sn[i] * fq[2 * i] + cs[i] * fq[2 * i + 1];
    ok[2 * i] = cs[i] * fk[2 * i] - sn[i] * fk[2 * i + 1];
    ok[2 * i + 1] = sn[i] * fk[2 * i] + cs[i] * fk[2 * i + 1];
  }
  *reinterpret_cast<uint4*>(q + dst) = pack8(oq);
  *reinterpret_cast<uint4*>(k + dst) = pack8(ok);
  *reinterpret_cast<uint4*>(v + dst) = uv;
}

}  // namespace ca

extern "C" ca_status ca_ln_modulate(const void* x, void* out, int B, int N, int D, const void* mod0,
                                    const void* mod1, int split, int64_t mod_stride,
                                    int shift_off, int scale_off, void* stream) {
  if (!out) return CA_ERR_INVALID_ARG;
  return ca::launch_ln<false>(const_cast<void*>(x), nullptr, 0, out, B, N, D, mod0, mod1, split,
                              mod_stride, 0, shift_off, scale_off, (cudaStream_t)stream);
}

extern "C" ca_status ca_gate_residual_ln_modulate(void* x, const void* y, int64_t y_stride,
                                                  void* out, int B, int N, int D, const void* mod0,
                                                  const void* mod1, int split, int64_t mod_stride,
                                                  int gate_off, int shift_off, int scale_off,
                                                  void* stream) {
  if (!y) return CA_ERR_INVALID_ARG;
  return ca::launch_ln<true>(x, y, y_stride, out, B, N, D, mod0, mod1, split, mod_stride, gate_off,
                             shift_off, scale_off, (cudaStream_t)stream);
}

extern "C" ca_status ca_qkv_prep(const void* qkv_a, int64_t row_stride_a, int Na, const void* qkv_b,
                                 int64_t row_stride_b, int Nb, const void* q_scale_a,
                                 const void* k_scale_a, const void* q_scale_b,
                                 const void* k_scale_b, const float* rope, int B, int H, void* q,
                                 void* k, void* v, void* q_pre, void* stream) {
  using namespace ca;
  if (Na < 0 || Nb < 0 || B < 0 || H <= 0 || !q || !k || !v || !rope) return CA_ERR_INVALID_ARG;
  if ((Na > 0 && (!qkv_a || !q_scale_a || !k_scale_a)) ||
      (Nb > 0 && (!qkv_b || !q_scale_b || !k_scale_b)))
    return CA_ERR_INVALID_ARG;
  if (row_stride_a % 8 || row_stride_b % 8) return CA_ERR_ALIGNMENT;
  const void* ptrs[] = {qkv_a, qkv_b, q_scale_a, k_scale_a, q_scale_b, k_scale_b, rope, q, k, v, q_pre};
  for (const void* p : ptrs) if (p && !aligned16(p)) return CA_ERR_ALIGNMENT;
  if (Na > 0 && row_stride_a < 3 * H * CA_HEAD_DIM) return CA_ERR_INVALID_ARG;
  if (Nb > 0 && row_stride_b < 3 * H * CA_HEAD_DIM) return CA_ERR_INVALID_ARG;
  const int64_t warps = (int64_t)B * (Na + Nb) * ((H + 1) / 2);
  if (warps == 0) return CA_OK;
  if (Na == 0) { qkv_a = qkv_b; q_scale_a = q_scale_b; k_scale_a = k_scale_b; }
  if (Nb == 0) { qkv_b = qkv_a; q_scale_b = q_scale_a; k_scale_b = k_scale_a; }
  const int64_t blocks = (warps + 7) / 8;
  if (blocks > 0x7fffffffLL) return CA_ERR_UNSUPPORTED;
  qkv_prep_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)qkv_a, row_stride_a, Na, (const __nv_bfloat16*)qkv_b, row_stride_b, Nb,
      (const __nv_bfloat16*)q_scale_a, (const __nv_bfloat16*)k_scale_a,
      (const __nv_bfloat16*)q_scale_b, (const __nv_bfloat16*)k_scale_b, rope, B, H,
      (__nv_bfloat16*)q, (__nv_bfloat16*)k, (__nv_bfloat16*)v, (__nv_bfloat16*)q_pre);
  CA_LAUNCH_CHECK();
  return CA_OK;
}
