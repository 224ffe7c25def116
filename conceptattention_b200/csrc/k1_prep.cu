// K1 family: the memory-bound elementwise/normalisation steps around the cuBLAS
// linears of a double-stream block.  All HBM-bound: one pass, 16-byte accesses,
// fp32 maths, bf16 storage, warp-shuffle reductions.
//
//   ca_ln_modulate                 modified_double_stream_block.py:88-89,94-95,100-101
//   ca_gate_residual_ln_modulate   modified_double_stream_block.py:194-202
//   ca_qkv_prep                    :91-92,97-98,103-104,110,125 + layers.py:63-84 + math.py:25-30
#include "ca_common.cuh"
#include "sm100_ptx.cuh"

#include <algorithm>

namespace ca {

using namespace ptx;

constexpr int kWarpsPerCta = 4;

// ---- row maths shared by both LN kernels (so they stay bit-identical) -----------------------
// A lane holds its share of a row as packed fp32 pairs (FADD2 / FFMA2: two elements per issue
// slot).  chunk = 8 consecutive elements = 4 pairs; lane i owns chunks i, i+32, ...
__device__ __forceinline__ void unpack8_pairs(const uint4& u, uint64_t* p) {
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int i = 0; i < 4; ++i)
    p[i] = pk64(__uint_as_float(w[i] << 16), __uint_as_float(w[i] & 0xffff0000u));
}
__device__ __forceinline__ uint32_t pack_pair(uint64_t p) {
  float lo, hi;
  upk64(p, lo, hi);
  return pack2(lo, hi);
}
// x' = bf16(x + gate*y) for one chunk, in place; returns the packed bf16 to store
__device__ __forceinline__ uint4 residual_chunk(uint64_t* v, const uint4& yraw, const uint4& graw) {
  uint64_t yp[4], gp[4];
  unpack8_pairs(yraw, yp);
  unpack8_pairs(graw, gp);
  uint32_t w[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    w[j] = pack_pair(fma2(gp[j], yp[j], v[j]));
    v[j] = pk64(__uint_as_float(w[j] << 16), __uint_as_float(w[j] & 0xffff0000u));
  }
  return make_uint4(w[0], w[1], w[2], w[3]);
}
template <int NCH>
__device__ __forceinline__ float row_sum(const uint64_t (&v)[NCH][4], int lane, int nchunks) {
  uint64_t s2 = pk64(0.f, 0.f);
#pragma unroll
  for (int k = 0; k < NCH; ++k)
    if (lane + k * 32 < nchunks) {
#pragma unroll
      for (int j = 0; j < 4; ++j) s2 = add2(s2, v[k][j]);
    }
  float a, b;
  upk64(s2, a, b);
  return warp_sum(a + b);
}
template <int NCH>
__device__ __forceinline__ float row_sqdev(const uint64_t (&v)[NCH][4], float mean, int lane, int nchunks) {
  const uint64_t nm2 = pk64(-mean, -mean);
  uint64_t s2 = pk64(0.f, 0.f);
#pragma unroll
  for (int k = 0; k < NCH; ++k)
    if (lane + k * 32 < nchunks) {
#pragma unroll
      for (int j = 0; j < 4; ++j) { const uint64_t d = add2(v[k][j], nm2); s2 = fma2(d, d, s2); }
    }
  float a, b;
  upk64(s2, a, b);
  return warp_sum(a + b);
}
// out = (1+scale) * ((v - mean) * rstd) + shift as two FMAs per pair:
//   t = v*rstd + (-mean*rstd);  o = t*a1 + sh      (a1 = 1+scale, sh = shift: fp32 pairs)
__device__ __forceinline__ uint4 norm_chunk(const uint64_t* v, uint64_t rstd2, uint64_t nmr2,
                                            const uint64_t* a1, const uint64_t* sh) {
  uint32_t w[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) w[j] = pack_pair(fma2(fma2(v[j], rstd2, nmr2), a1[j], sh[j]));
  return make_uint4(w[0], w[1], w[2], w[3]);
}
// a1 / sh pairs of one chunk straight from the bf16 modulation row in global memory
__device__ __forceinline__ void mod_chunk_global(const __nv_bfloat16* mod, int shift_off, int scale_off,
                                                 int c, uint64_t* a1, uint64_t* sh) {
  unpack8_pairs(*reinterpret_cast<const uint4*>(mod + shift_off + c * 8), sh);
  unpack8_pairs(*reinterpret_cast<const uint4*>(mod + scale_off + c * 8), a1);
  const uint64_t one2 = pk64(1.f, 1.f);
#pragma unroll
  for (int j = 0; j < 4; ++j) a1[j] = add2(a1[j], one2);
}

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

  uint64_t v[NCH][4];
#pragma unroll
  for (int i = 0; i < NCH; ++i) {
    const int c = lane + i * 32;
    if (c < nchunks) {
      const uint4 u = kResidual ? *reinterpret_cast<const uint4*>(xr + c * 8) : ld_stream16(xr + c * 8);
      unpack8_pairs(u, v[i]);
      if (kResidual) {
        *reinterpret_cast<uint4*>(xr + c * 8) = residual_chunk(
            v[i], ld_stream16(y + (int64_t)row * y_stride + c * 8),
            *reinterpret_cast<const uint4*>(mod + gate_off + c * 8));
      }
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j) v[i][j] = pk64(0.f, 0.f);
    }
  }
  if (out == nullptr) return;
  const float mean = row_sum<NCH>(v, lane, nchunks) / D;
  const float rstd = rsqrtf(row_sqdev<NCH>(v, mean, lane, nchunks) / D + 1e-6f);
  const uint64_t rstd2 = pk64(rstd, rstd), nmr2 = pk64(-mean * rstd, -mean * rstd);
#pragma unroll
  for (int i = 0; i < NCH; ++i) {
    const int c = lane + i * 32;
    if (c < nchunks) {
      uint64_t a1[4], sh[4];
      mod_chunk_global(mod, shift_off, scale_off, c, a1, sh);
      *reinterpret_cast<uint4*>(out + (int64_t)row * D + c * 8) = norm_chunk(v[i], rstd2, nmr2, a1, sh);
    }
  }
}

// Streaming form of the same row-wise op for the large token streams.  Persistent: one CTA of
// 16 warps per SM, CTA c owns the contiguous row range [c*rows/G, (c+1)*rows/G), its warps take
// those rows round-robin.  Rows are staged through shared memory by the TMA engine
// (cp.async.bulk 1-D, one mbarrier per slot): a warp copies its row smem -> registers, re-arms
// the slot for its next row at once and computes while that load is in flight, so the bytes in
// flight per SM (96-192 KB) do not depend on occupancy and the grid is exactly the SM count (no
// wave tail).  When every row of the CTA uses the same modulation row (always true for the image
// and single streams at B = 1), (1+scale) and shift are converted to fp32 once per CTA into
// shared memory instead of once per row.  Same maths as ln_modulate_kernel: bit-identical.
constexpr int kStreamWarps = 16;

template <int NCH, bool kResidual>
__global__ void __launch_bounds__(kStreamWarps * 32, 1)
ln_stream_kernel(__nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ y,
                 int64_t y_stride, __nv_bfloat16* __restrict__ out, int rows, int N, int D,
                 const __nv_bfloat16* __restrict__ mod0, const __nv_bfloat16* __restrict__ mod1,
                 int split, int64_t mod_stride, int gate_off, int shift_off, int scale_off,
                 int slots) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t row_bytes = (uint32_t)D * 2u;
  const uint32_t slot_bytes = row_bytes * (kResidual ? 2u : 1u);
  // layout: [a1 fp32 D | sh fp32 D] (only with an output) | [gate bf16 D] (residual forms) |
  //         row slots | mbarriers
  float* a1_s = reinterpret_cast<float*>(smem);
  float* sh_s = a1_s + D;
  uint8_t* gate_s = smem + (out != nullptr ? (size_t)D * 8 : 0);
  uint8_t* slots_base = gate_s + (kResidual ? (size_t)D * 2 : 0);
  uint8_t* wbase = slots_base + (size_t)warp * slots * slot_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(slots_base + (size_t)kStreamWarps * slots * slot_bytes) + warp * slots;
  const int64_t r_begin = (int64_t)blockIdx.x * rows / gridDim.x;
  const int64_t r_end = (int64_t)(blockIdx.x + 1) * rows / gridDim.x;
  const int my_n = (int)((r_end - r_begin - warp + kStreamWarps - 1) / kStreamWarps);   // may be <= 0
  const int nchunks = D >> 3;

  auto issue = [&](int i) {                               // lane 0 only
    const int slot = i % slots;
    const int64_t row = r_begin + warp + (int64_t)kStreamWarps * i;
    uint8_t* dst = wbase + (size_t)slot * slot_bytes;
    mbar_expect_tx(&bars[slot], slot_bytes);
    bulk_load_1d(dst, x + row * D, row_bytes, &bars[slot]);
    if (kResidual) bulk_load_1d(dst + row_bytes, y + row * y_stride, row_bytes, &bars[slot]);
  };
  if (lane == 0) {
    for (int s = 0; s < slots; ++s) mbar_init(&bars[s], 1);
    fence_barrier_init();
    fence_proxy_async_smem();
    for (int i = 0; i < slots && i < my_n; ++i) issue(i);
  }
  // one modulation row for the whole CTA?  (first and last row in the same batch element and
  // on the same side of `split`)
  bool cached = false;
  if (r_end > r_begin) {
    const int b0 = (int)(r_begin / N), b1 = (int)((r_end - 1) / N);
    const int n0 = (int)(r_begin - (int64_t)b0 * N), n1 = (int)(r_end - 1 - (int64_t)b1 * N);
    cached = b0 == b1 && (n0 < split) == (n1 < split);
    if (cached) {
      const __nv_bfloat16* mod = (n0 < split ? mod0 : mod1) + (int64_t)b0 * mod_stride;
      for (int c = threadIdx.x; c < nchunks; c += blockDim.x) {
        if (kResidual)
          reinterpret_cast<uint4*>(gate_s)[c] = *reinterpret_cast<const uint4*>(mod + gate_off + c * 8);
        if (out != nullptr) {
          uint64_t a1[4], sh[4];
          mod_chunk_global(mod, shift_off, scale_off, c, a1, sh);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            reinterpret_cast<uint64_t*>(a1_s)[c * 4 + j] = a1[j];
            reinterpret_cast<uint64_t*>(sh_s)[c * 4 + j] = sh[j];
          }
        }
      }
    }
  }
  __syncthreads();                                        // cache + this warp's barriers visible

  for (int i = 0; i < my_n; ++i) {
    const int slot = i % slots;
    const int64_t row = r_begin + warp + (int64_t)kStreamWarps * i;
    const int b = (int)(row / N), n = (int)(row - (int64_t)b * N);
    const __nv_bfloat16* mod = (n < split ? mod0 : mod1) + (int64_t)b * mod_stride;
    const uint8_t* src = wbase + (size_t)slot * slot_bytes;
    mbar_wait(&bars[slot], (i / slots) & 1);

    uint64_t v[NCH][4];
#pragma unroll
    for (int k = 0; k < NCH; ++k) {
      const int c = lane + k * 32;
      if (c < nchunks) {
        unpack8_pairs(*reinterpret_cast<const uint4*>(src + c * 16), v[k]);
        if (kResidual) {
          const uint4 graw = cached ? reinterpret_cast<const uint4*>(gate_s)[c]
                                    : *reinterpret_cast<const uint4*>(mod + gate_off + c * 8);
          *reinterpret_cast<uint4*>(x + row * D + c * 8) = residual_chunk(
              v[k], *reinterpret_cast<const uint4*>(src + row_bytes + c * 16), graw);
        }
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) v[k][j] = pk64(0.f, 0.f);
      }
    }
    __syncwarp();                                         // every lane has read the slot
    if (lane == 0 && i + slots < my_n) { fence_proxy_async_smem(); issue(i + slots); }
    if (out == nullptr) continue;

    const float mean = row_sum<NCH>(v, lane, nchunks) / D;
    const float rstd = rsqrtf(row_sqdev<NCH>(v, mean, lane, nchunks) / D + 1e-6f);
    const uint64_t rstd2 = pk64(rstd, rstd), nmr2 = pk64(-mean * rstd, -mean * rstd);
#pragma unroll
    for (int k = 0; k < NCH; ++k) {
      const int c = lane + k * 32;
      if (c < nchunks) {
        uint64_t a1[4], sh[4];
        if (cached) {
          const ulonglong2 a_lo = *reinterpret_cast<const ulonglong2*>(a1_s + c * 8);
          const ulonglong2 a_hi = *reinterpret_cast<const ulonglong2*>(a1_s + c * 8 + 4);
          const ulonglong2 s_lo = *reinterpret_cast<const ulonglong2*>(sh_s + c * 8);
          const ulonglong2 s_hi = *reinterpret_cast<const ulonglong2*>(sh_s + c * 8 + 4);
          a1[0] = a_lo.x; a1[1] = a_lo.y; a1[2] = a_hi.x; a1[3] = a_hi.y;
          sh[0] = s_lo.x; sh[1] = s_lo.y; sh[2] = s_hi.x; sh[3] = s_hi.y;
        } else {
          mod_chunk_global(mod, shift_off, scale_off, c, a1, sh);
        }
        *reinterpret_cast<uint4*>(out + row * D + c * 8) = norm_chunk(v[k], rstd2, nmr2, a1, sh);
      }
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
  // Large token streams: persistent TMA-staged kernel, grid = SM count.  Small ones (the
  // text+concept rows) and rows too wide for the staging buffers keep the one-warp-per-row form.
  const size_t slot_bytes = (size_t)D * 2 * (kResidual ? 2 : 1);
  // fp32 (1+scale) and shift (forms with an output) + bf16 gate (residual forms)
  const size_t cache_bytes = (out ? (size_t)D * 8 : 0) + (kResidual ? (size_t)D * 2 : 0);
  const size_t budget = 226u * 1024u - cache_bytes;                // 227 KB per CTA on sm_100
  const int slots = (int)std::min<size_t>(2, budget / (kStreamWarps * (slot_bytes + 8) + 256));
  // Measured on B200 (scripts/bench_k1.py): the plain LN (one input stream) is faster in the
  // one-warp-per-row form (0.70 vs 0.60 of HBM at 16 K rows); the residual forms (two input
  // streams, in-place update) are faster staged (0.74 / 0.79 vs 0.65 / 0.61).
  const bool stream = kResidual;
  if (stream && rows >= 1024 && slots >= 1 && nch <= 16) {
    static int sm_count = 0;
    if (sm_count == 0) {
      int dev = 0;
      CA_CUDA_OK(cudaGetDevice(&dev));
      CA_CUDA_OK(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
    }
    const int grid_s = std::min(sm_count, (rows + kStreamWarps - 1) / kStreamWarps);
    const int smem = (int)(cache_bytes + kStreamWarps * slots * (slot_bytes + 8) + 256);
#define CA_LNS_CASE(NCH)                                                                          \
  do {                                                                                            \
    auto kern = ln_stream_kernel<NCH, kResidual>;                                                 \
    CA_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));    \
    kern<<<grid_s, kStreamWarps * 32, smem, st>>>(                                                \
        (__nv_bfloat16*)x, (const __nv_bfloat16*)y, y_stride, (__nv_bfloat16*)out, rows, N, D,    \
        (const __nv_bfloat16*)mod0, (const __nv_bfloat16*)mod1, split, mod_stride, gate_off,      \
        shift_off, scale_off, slots);                                                             \
  } while (0)
    if (nch <= 1) CA_LNS_CASE(1);
    else if (nch <= 2) CA_LNS_CASE(2);
    else if (nch <= 4) CA_LNS_CASE(4);
    else if (nch <= 8) CA_LNS_CASE(8);
    else if (nch <= 12) CA_LNS_CASE(12);
    else CA_LNS_CASE(16);
#undef CA_LNS_CASE
    CA_LAUNCH_CHECK();
    return CA_OK;
  }
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
                __nv_bfloat16* __restrict__ v, __nv_bfloat16* __restrict__ q_pre,
                int group_heads, int64_t group_stride) {
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
  // grouped form (sequence-parallel exchange): heads [g*G, (g+1)*G) of q, k and v go to group g's
  // slab, so each destination rank's chunk of the all-to-all send buffer is contiguous
  const int64_t dst_g = group_heads
      ? (int64_t)(h / group_heads) * group_stride + ((int64_t)(h % group_heads) * Ntot + n) * CA_HEAD_DIM + ch * 8
      : dst;
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
  *reinterpret_cast<uint4*>(q + dst_g) = pack8(oq);
  *reinterpret_cast<uint4*>(k + dst_g) = pack8(ok);
  *reinterpret_cast<uint4*>(v + dst_g) = uv;
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

extern "C" ca_status ca_qkv_prep_grouped(const void* qkv_a, int64_t row_stride_a, int Na,
                                         const void* qkv_b, int64_t row_stride_b, int Nb,
                                         const void* q_scale_a, const void* k_scale_a,
                                         const void* q_scale_b, const void* k_scale_b,
                                         const float* rope, int B, int H, void* q, void* k, void* v,
                                         void* q_pre, int group_heads, int64_t group_stride,
                                         void* stream) {
  using namespace ca;
  if (group_heads < 0 || (group_heads > 0 && (B != 1 || H % group_heads || group_stride % 8 ||
                                              group_stride < (int64_t)group_heads * (Na + Nb) * CA_HEAD_DIM)))
    return CA_ERR_INVALID_ARG;
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
      (__nv_bfloat16*)q, (__nv_bfloat16*)k, (__nv_bfloat16*)v, (__nv_bfloat16*)q_pre, group_heads,
      group_stride);
  CA_LAUNCH_CHECK();
  return CA_OK;
}

extern "C" ca_status ca_qkv_prep(const void* qkv_a, int64_t row_stride_a, int Na, const void* qkv_b,
                                 int64_t row_stride_b, int Nb, const void* q_scale_a,
                                 const void* k_scale_a, const void* q_scale_b,
                                 const void* k_scale_b, const float* rope, int B, int H, void* q,
                                 void* k, void* v, void* q_pre, void* stream) {
  return ca_qkv_prep_grouped(qkv_a, row_stride_a, Na, qkv_b, row_stride_b, Nb, q_scale_a, k_scale_a,
                             q_scale_b, k_scale_b, rope, B, H, q, k, v, q_pre, 0, 0, stream);
}
