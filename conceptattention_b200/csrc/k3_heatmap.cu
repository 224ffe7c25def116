// K3 + K4: heatmap logits (image . concept), normalisation over the concept axis
// and accumulation over the selected (timestep, layer) pairs.
//
// Restates concept_attention/concept_attention_pipeline.py:43-82 as ONE pass over
// the image vectors: HBM-bound (AI ~ C FLOP/B), so the design is about bytes:
//  - the [C,K] concept matrix is staged once per CTA in shared memory;
//  - a warp owns R image rows and streams them with 16-byte no-allocate loads,
//    lane i taking chunks i, i+32, ... so every request is a full 512 B line set;
//  - fp32 partial sums are warp-shuffle reduced, then one lane per row applies
//    softmax / sparsemax / entmax15 in registers and updates the fp32 accumulator
//    acc[b,c,p] (coalesced across the R rows of neighbouring warps per concept).
// The 4 x [Nt,L,B,...] stacks of the reference (a8 in SURVEY.md) never exist.
#include "ca_common.cuh"
#include "heatmap_norm.cuh"

namespace ca {

constexpr int kHmWarps = 8;

template <int R, int CT, bool kLogitsOnly>
__global__ void __launch_bounds__(kHmWarps * 32)
heatmap_kernel(const HmArgs a) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  __nv_bfloat16* scon = reinterpret_cast<__nv_bfloat16*>(smem_raw);   // [C][K]
  const int K = a.nseg * a.seg_len, nchunks = K >> 3;
  const int b = blockIdx.y;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

  for (int i = threadIdx.x; i < a.C * nchunks; i += blockDim.x) {
    const int c = i / nchunks, ch = i - c * nchunks;
    const int e = ch * 8, seg = e / a.seg_len, off = e - seg * a.seg_len;
    *reinterpret_cast<uint4*>(scon + (int64_t)c * K + e) = *reinterpret_cast<const uint4*>(
        a.con + b * a.csb + seg * a.css + c * a.csc + off);
  }
  __syncthreads();

  const int row0 = (blockIdx.x * kHmWarps + warp) * R;
  if (row0 >= a.P) return;
  const __nv_bfloat16* base = a.img + b * a.isb;
  int64_t roff[R];
#pragma unroll
  for (int r = 0; r < R; ++r) roff[r] = (int64_t)min(row0 + r, a.P - 1) * a.isp;

  float sum[R][CT];
#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int c = 0; c < CT; ++c) sum[r][c] = 0.f;

  for (int ch = lane; ch < nchunks; ch += 32) {
    const int e = ch * 8, seg = e / a.seg_len, off = e - seg * a.seg_len;
    const int64_t so = seg * a.iss + off;
    uint4 u[R];
#pragma unroll
    for (int r = 0; r < R; ++r) u[r] = ld_stream16(base + roff[r] + so);
    float x[R][8];
#pragma unroll
    for (int r = 0; r < R; ++r) unpack8(u[r], x[r]);
#pragma unroll
    for (int c = 0; c < CT; ++c) {
      if (c < a.C) {
        float w[8];
        unpack8(*reinterpret_cast<const uint4*>(scon + (int64_t)c * K + e), w);
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
          for (int j = 0; j < 8; ++j) sum[r][c] = fmaf(x[r][j], w[j], sum[r][c]);
      }
    }
  }
#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int c = 0; c < CT; ++c) sum[r][c] = warp_sum(sum[r][c]);

#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int p = row0 + r;
    if (lane == r && p < a.P) {
      float* dst = a.acc + (int64_t)b * a.C * a.P + p;
      if (kLogitsOnly) {
#pragma unroll
        for (int c = 0; c < CT; ++c) if (c < a.C) dst[(int64_t)c * a.P] = sum[r][c];
      } else {
        normalise<CT>(sum[r], a.C, a.norm, a.round_bf16 != 0);
        float old[CT];                       // all loads before the first store (no serialised RMW)
#pragma unroll
        for (int c = 0; c < CT; ++c) old[c] = c < a.C ? __ldcg(dst + (int64_t)c * a.P) : 0.f;
#pragma unroll
        for (int c = 0; c < CT; ++c) if (c < a.C) dst[(int64_t)c * a.P] = old[c] + sum[r][c];
        if (float* pb = pair_base(a)) {
          float* pd = pb + (int64_t)b * a.C * a.P + p;
#pragma unroll
          for (int c = 0; c < CT; ++c) if (c < a.C) pd[(int64_t)c * a.P] = sum[r][c];
        }
      }
    }
  }
}

// K4 alone: one thread per (b,p).
template <int CT>
__global__ void norm_accumulate_kernel(const float* __restrict__ logits, float* __restrict__ acc,
                                       int B, int P, int C, int norm, int round_bf16,
                                       float* __restrict__ pair) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)B * P) return;
  const int b = (int)(i / P), p = (int)(i - (int64_t)b * P);
  float z[CT];
#pragma unroll
  for (int c = 0; c < CT; ++c) z[c] = c < C ? logits[((int64_t)b * C + c) * P + p] : 0.f;
  float old[CT];
#pragma unroll
  for (int c = 0; c < CT; ++c) old[c] = c < C ? __ldcg(acc + ((int64_t)b * C + c) * P + p) : 0.f;
  normalise<CT>(z, C, norm, round_bf16 != 0);
#pragma unroll
  for (int c = 0; c < CT; ++c) if (c < C) acc[((int64_t)b * C + c) * P + p] = old[c] + z[c];
  if (pair) {
#pragma unroll
    for (int c = 0; c < CT; ++c) if (c < C) pair[((int64_t)b * C + c) * P + p] = z[c];
  }
}

// Column-wise min-shift / sum normalisation of a [C,K] bf16 matrix (utils.py:35-44).
__global__ void concept_linear_normalize_kernel(const __nv_bfloat16* __restrict__ in,
                                                __nv_bfloat16* __restrict__ out, int C, int64_t K) {
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= K) return;
  float mn = INFINITY;
  for (int c = 0; c < C; ++c) mn = fminf(mn, __bfloat162float(in[c * K + k]));
  float s = 0.f;
  for (int c = 0; c < C; ++c) s += bf16_round(__bfloat162float(in[c * K + k]) - mn);
  s = bf16_round(s);
  if (s == 0.f) s = 1.f;
  for (int c = 0; c < C; ++c)
    out[c * K + k] = __float2bfloat16_rn(bf16_round(__bfloat162float(in[c * K + k]) - mn) / s);
}


template <bool kLogitsOnly>
static ca_status launch_heatmap(const HmArgs& a, cudaStream_t st) {
  if (!a.img || !a.con || !a.acc || a.B < 0 || a.P < 0 || a.C <= 0 || a.nseg <= 0 || a.seg_len <= 0)
    return CA_ERR_INVALID_ARG;
  if (a.C > CA_MAX_CONCEPTS || a.seg_len % 8) return CA_ERR_UNSUPPORTED;
  if (a.norm < CA_NORM_SOFTMAX || a.norm > CA_NORM_NONE) return CA_ERR_INVALID_ARG;
  if (!aligned16(a.img) || !aligned16(a.con) || (a.isb | a.iss | a.isp | a.csb | a.css | a.csc) % 8)
    return CA_ERR_ALIGNMENT;
  if (a.B == 0 || a.P == 0) return CA_OK;
  {                                   // TMA/tcgen05 streaming variant when the shape fits it
    const ca_status s = launch_heatmap_tc(a, kLogitsOnly, st);   // (seg_len % 64 == 0); else SIMT below
    if (s != CA_ERR_UNSUPPORTED) return s;
  }
  const size_t smem = (size_t)a.C * a.nseg * a.seg_len * 2;
  if (smem > 220 * 1024) return CA_ERR_UNSUPPORTED;
#define CA_HM_LAUNCH(R, CT)                                                                     \
  do {                                                                                          \
    auto kern = heatmap_kernel<R, CT, kLogitsOnly>;                                             \
    if (smem > 48 * 1024)                                                                       \
      CA_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    dim3 grid((a.P + kHmWarps * R - 1) / (kHmWarps * R), a.B);                                  \
    kern<<<grid, kHmWarps * 32, smem, st>>>(a);                                                 \
  } while (0)
  if (a.C <= 4) CA_HM_LAUNCH(2, 4);
  else if (a.C <= 8) CA_HM_LAUNCH(2, 8);
  else if (a.C <= 16) CA_HM_LAUNCH(2, 16);
  else CA_HM_LAUNCH(1, 32);
#undef CA_HM_LAUNCH
  CA_LAUNCH_CHECK();
  return CA_OK;
}

}  // namespace ca


extern "C" ca_status ca_heatmap_accumulate(const void* img, int64_t isb, int64_t iss, int64_t isp,
                                           const void* con, int64_t csb, int64_t css, int64_t csc,
                                           int nseg, int seg_len, int B, int P, int C, int norm,
                                           int round_bf16, float* acc, void* stream) {
  ca::HmArgs a{(const __nv_bfloat16*)img, isb, iss, isp, (const __nv_bfloat16*)con, csb, css, csc,
               nseg, seg_len, B, P, C, norm, round_bf16, acc};
  return ca::launch_heatmap<false>(a, (cudaStream_t)stream);
}

extern "C" ca_status ca_heatmap_accumulate_pair(const void* img, int64_t isb, int64_t iss, int64_t isp,
                                                const void* con, int64_t csb, int64_t css, int64_t csc,
                                                int nseg, int seg_len, int B, int P, int C, int norm,
                                                int round_bf16, float* acc, float* pair,
                                                int64_t pair_stride, const int* pair_slot,
                                                int pair_slot_off, void* stream) {
  ca::HmArgs a{(const __nv_bfloat16*)img, isb, iss, isp, (const __nv_bfloat16*)con, csb, css, csc,
               nseg, seg_len, B, P, C, norm, round_bf16, acc};
  a.pair = pair; a.pair_stride = pair_stride; a.pair_slot = pair_slot; a.pair_slot_off = pair_slot_off;
  if (pair && pair_stride < (int64_t)B * C * P) return CA_ERR_INVALID_ARG;
  return ca::launch_heatmap<false>(a, (cudaStream_t)stream);
}

// Output-space and cross-attention-space maps of one layer in ONE launch (gridDim.z = 2 of the
// TMA/tcgen05 kernel) when both fit it; otherwise the two launches of ca_heatmap_accumulate_pair.
extern "C" ca_status ca_heatmap_accumulate_both(
    const void* img0, int64_t isb0, int64_t iss0, int64_t isp0, const void* con0, int64_t csb0, int64_t css0,
    int64_t csc0, int nseg0, int seg_len0, float* acc0, int pair_slot_off0,
    const void* img1, int64_t isb1, int64_t iss1, int64_t isp1, const void* con1, int64_t csb1, int64_t css1,
    int64_t csc1, int nseg1, int seg_len1, float* acc1, int pair_slot_off1,
    int B, int P, int C, int norm, int round_bf16, float* pair, int64_t pair_stride, const int* pair_slot,
    void* stream) {
  ca::HmArgs a{(const __nv_bfloat16*)img0, isb0, iss0, isp0, (const __nv_bfloat16*)con0, csb0, css0, csc0,
               nseg0, seg_len0, B, P, C, norm, round_bf16, acc0};
  ca::HmArgs b{(const __nv_bfloat16*)img1, isb1, iss1, isp1, (const __nv_bfloat16*)con1, csb1, css1, csc1,
               nseg1, seg_len1, B, P, C, norm, round_bf16, acc1};
  a.pair = b.pair = pair; a.pair_stride = b.pair_stride = pair_stride; a.pair_slot = b.pair_slot = pair_slot;
  a.pair_slot_off = pair_slot_off0; b.pair_slot_off = pair_slot_off1;
  if (pair && pair_stride < (int64_t)B * C * P) return CA_ERR_INVALID_ARG;
  const ca::HmArgs* both[2] = {&a, &b};
  bool fits = true;
  for (const ca::HmArgs* x : both) {
    if (!x->img || !x->con || !x->acc || B < 0 || P < 0 || C <= 0 || x->nseg <= 0 || x->seg_len <= 0)
      return CA_ERR_INVALID_ARG;
    if (C > CA_MAX_CONCEPTS || x->seg_len % 8) return CA_ERR_UNSUPPORTED;
    if (!ca::aligned16(x->img) || !ca::aligned16(x->con) ||
        (x->isb | x->iss | x->isp | x->csb | x->css | x->csc) % 8)
      return CA_ERR_ALIGNMENT;
  }
  if (norm < CA_NORM_SOFTMAX || norm > CA_NORM_NONE) return CA_ERR_INVALID_ARG;
  if (B == 0 || P == 0) return CA_OK;
  if (fits) {
    const ca_status s = ca::launch_heatmap_tc_dual(a, b, (cudaStream_t)stream);
    if (s != CA_ERR_UNSUPPORTED) return s;
  }
  ca_status s = ca::launch_heatmap<false>(a, (cudaStream_t)stream);
  if (s != CA_OK) return s;
  return ca::launch_heatmap<false>(b, (cudaStream_t)stream);
}

extern "C" ca_status ca_heatmap_logits(const void* img, int64_t isb, int64_t iss, int64_t isp,
                                       const void* con, int64_t csb, int64_t css, int64_t csc,
                                       int nseg, int seg_len, int B, int P, int C, float* logits,
                                       void* stream) {
  ca::HmArgs a{(const __nv_bfloat16*)img, isb, iss, isp, (const __nv_bfloat16*)con, csb, css, csc,
               nseg, seg_len, B, P, C, CA_NORM_NONE, 0, logits};
  return ca::launch_heatmap<true>(a, (cudaStream_t)stream);
}

extern "C" ca_status ca_norm_accumulate_pair(const float* logits, int B, int P, int C, int norm,
                                             int round_bf16, float* acc, float* pair, void* stream) {
  if (!logits || !acc || B < 0 || P < 0 || C <= 0) return CA_ERR_INVALID_ARG;
  if (C > CA_MAX_CONCEPTS) return CA_ERR_UNSUPPORTED;
  if (norm < CA_NORM_SOFTMAX || norm > CA_NORM_NONE) return CA_ERR_INVALID_ARG;
  const int64_t n = (int64_t)B * P;
  if (n == 0) return CA_OK;
  const unsigned blocks = (unsigned)((n + 255) / 256);
  auto st = (cudaStream_t)stream;
  if (C <= 4) ca::norm_accumulate_kernel<4><<<blocks, 256, 0, st>>>(logits, acc, B, P, C, norm, round_bf16, pair);
  else if (C <= 8) ca::norm_accumulate_kernel<8><<<blocks, 256, 0, st>>>(logits, acc, B, P, C, norm, round_bf16, pair);
  else if (C <= 16) ca::norm_accumulate_kernel<16><<<blocks, 256, 0, st>>>(logits, acc, B, P, C, norm, round_bf16, pair);
  else ca::norm_accumulate_kernel<32><<<blocks, 256, 0, st>>>(logits, acc, B, P, C, norm, round_bf16, pair);
  CA_LAUNCH_CHECK();
  return CA_OK;
}

extern "C" ca_status ca_norm_accumulate(const float* logits, int B, int P, int C, int norm,
                                        int round_bf16, float* acc, void* stream) {
  return ca_norm_accumulate_pair(logits, B, P, C, norm, round_bf16, acc, nullptr, stream);
}

extern "C" ca_status ca_concept_linear_normalize(const void* con, void* out, int C, int64_t K,
                                                 void* stream) {
  if (!con || !out || C <= 0 || K <= 0) return CA_ERR_INVALID_ARG;
  ca::concept_linear_normalize_kernel<<<(unsigned)((K + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)con, (__nv_bfloat16*)out, C, K);
  CA_LAUNCH_CHECK();
  return CA_OK;
}
