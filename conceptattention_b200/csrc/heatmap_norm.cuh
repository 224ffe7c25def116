// Shared by the SIMT (k3_heatmap.cu) and tcgen05 (k3_heatmap_tc.cu) heatmap kernels:
// the normalisers over the concept axis (concept_attention_pipeline.py:64-71) and the
// argument block.
#pragma once
#include "ca_common.cuh"

namespace ca {

template <int CT>
__device__ __forceinline__ void normalise(float (&z)[CT], int C, int norm, bool round_bf16) {
  if (round_bf16) {
#pragma unroll
    for (int c = 0; c < CT; ++c) z[c] = bf16_round(z[c]);
  }
  if (norm == CA_NORM_NONE) return;
  float m = -INFINITY;
#pragma unroll
  for (int c = 0; c < CT; ++c) if (c < C) m = fmaxf(m, z[c]);
  if (norm == CA_NORM_SOFTMAX) {
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < CT; ++c) { z[c] = (c < C) ? __expf(z[c] - m) : 0.f; s += z[c]; }
    const float inv = 1.f / s;
#pragma unroll
    for (int c = 0; c < CT; ++c) z[c] *= inv;
  } else {
    // Sort-free O(C^2) threshold search (C <= 32).  For element i let
    // k_i = #{z_j >= z_i}, s_i = sum of those, q_i = sum of their squares.
    const bool ent = (norm == CA_NORM_ENTMAX15);
    if (ent) {
#pragma unroll
      for (int c = 0; c < CT; ++c) z[c] = (z[c] - m) * 0.5f;
    }
    float tau = 0.f, best_k = 0.f;
#pragma unroll
    for (int i = 0; i < CT; ++i) {
      if (i < C) {
        float k = 0.f, s = 0.f, q = 0.f;
#pragma unroll
        for (int j = 0; j < CT; ++j) {
          if (j < C && z[j] >= z[i]) { k += 1.f; s += z[j]; q += z[j] * z[j]; }
        }
        float t;
        bool in_support;
        if (!ent) {                       // sparsemax: 1 + k z_i > s
          t = (s - 1.f) / k;
          in_support = 1.f + k * z[i] > s;
        } else {                          // entmax-1.5 (Peters et al. 2019, Alg. 2)
          const float mean = s / k, msq = q / k;
          const float delta = fmaxf((1.f - k * (msq - mean * mean)) / k, 0.f);
          t = mean - sqrtf(delta);
          in_support = t <= z[i];
        }
        if (in_support && k > best_k) { best_k = k; tau = t; }
      }
    }
#pragma unroll
    for (int c = 0; c < CT; ++c) {
      const float d = (c < C) ? fmaxf(z[c] - tau, 0.f) : 0.f;
      z[c] = ent ? d * d : d;
    }
  }
  if (round_bf16) {
#pragma unroll
    for (int c = 0; c < CT; ++c) z[c] = bf16_round(z[c]);
  }
}

struct HmArgs {
  const __nv_bfloat16* img; int64_t isb, iss, isp;
  const __nv_bfloat16* con; int64_t csb, css, csc;
  int nseg, seg_len, B, P, C, norm, round_bf16;
  float* acc;      // fused: acc += norm(logits);   logits-only: acc = logits
  // optional per-(timestep, layer) slot: pair[slot][b,c,p] = norm(logits) in the same pass
  // (concept_attention_pipeline.py:76-77 before the mean).  slot = *pair_slot (a DEVICE int,
  // so a captured forward can be replayed for every timestep) + pair_slot_off.
  float* pair = nullptr; int64_t pair_stride = 0; const int* pair_slot = nullptr; int pair_slot_off = 0;
};

__device__ __forceinline__ float* pair_base(const HmArgs& a) {
  if (!a.pair) return nullptr;
  const int slot = (a.pair_slot ? __ldg(a.pair_slot) : 0) + a.pair_slot_off;
  return a.pair + (int64_t)slot * a.pair_stride;
}


// tcgen05/TMA variant (k3_heatmap_tc.cu).  Returns CA_ERR_UNSUPPORTED when the shape does not
// fit it (seg_len % 64 != 0, ...): the caller then uses the SIMT kernel.
ca_status launch_heatmap_tc(const HmArgs& a, bool logits_only, cudaStream_t st);
ca_status launch_heatmap_tc_dual(const HmArgs& a, const HmArgs& b, cudaStream_t st);

}  // namespace ca
