// K5: finalise the accumulated heatmaps and turn them into masks on the device
// (the reference moves them to the host first, segmentation.py:77).
//   ca_finalize_heatmaps  mean over the selected (t,l)   concept_attention_pipeline.py:78-82
//   ca_gaussian_blur3     segmentation.py:57-58 (torchvision gaussian_blur k=3, sigma=1, reflect)
//   ca_argmax_mask        experiments/pascal_voc_segmentation/multi_class_segmentation.py:77
//                         + run_multi_class_seg_experiment.py:25-31
//   ca_binary_masks       segmentation.py:60-61
#include "ca_common.cuh"

namespace ca {

__global__ void finalize_kernel(const float* __restrict__ acc, float* __restrict__ heat, int64_t n,
                                float inv_count, int round_bf16) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float v = acc[i] * inv_count;
  heat[i] = round_bf16 ? bf16_round(v) : v;
}

__global__ void blur3_kernel(const float* __restrict__ in, float* __restrict__ out, int M, int h,
                             int w) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)M * h * w) return;
  const int x = (int)(i % w), y = (int)((i / w) % h);
  const float* src = in + (i / ((int64_t)h * w)) * h * w;
  // 1-D taps exp(-x^2/2) on {-1,0,1}, normalised
  const float e = 0.60653065971263342f, s = 1.f / (1.f + 2.f * e);
  const float k[3] = {e * s, s, e * s};
  float accv = 0.f;
#pragma unroll
  for (int dy = -1; dy <= 1; ++dy) {
    int yy = y + dy;
    yy = yy < 0 ? -yy : (yy >= h ? 2 * h - 2 - yy : yy);     // reflect (no edge repeat)
#pragma unroll
    for (int dx = -1; dx <= 1; ++dx) {
      int xx = x + dx;
      xx = xx < 0 ? -xx : (xx >= w ? 2 * w - 2 - xx : xx);
      accv += k[dy + 1] * k[dx + 1] * src[(int64_t)yy * w + xx];
    }
  }
  out[i] = accv;
}

__global__ void argmax_kernel(const float* __restrict__ heat, int64_t* __restrict__ mask, int B,
                              int C, int P, int num_background) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)B * P) return;
  const int b = (int)(i / P), p = (int)(i - (int64_t)b * P);
  const float* src = heat + (int64_t)b * C * P + p;
  float best = src[0];
  int arg = 0;
  for (int c = 1; c < C; ++c) {
    const float v = src[(int64_t)c * P];
    if (v > best || (v != v && best == best)) { best = v; arg = c; }   // first max; NaN wins like torch
  }
  mask[i] = arg < num_background ? 0 : arg;
}

// One CTA per map: block-reduce the mean in fp32 pairwise order, then compare.
__global__ void __launch_bounds__(256)
binary_mask_kernel(const float* __restrict__ heat, uint8_t* __restrict__ mask, int P) {
  __shared__ float part[8];
  __shared__ float mean_s;
  const float* src = heat + (int64_t)blockIdx.x * P;
  float s = 0.f;
  for (int i = threadIdx.x; i < P; i += 256) s += src[i];
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < 8; ++i) t += part[i];
    mean_s = t / P;
  }
  __syncthreads();
  const float m = mean_s;
  for (int i = threadIdx.x; i < P; i += 256) mask[(int64_t)blockIdx.x * P + i] = src[i] > m;
}

}  // namespace ca

extern "C" ca_status ca_finalize_heatmaps(const float* acc, float* heat, int64_t n, float inv_count,
                                          int round_bf16, void* stream) {
  if (!acc || !heat || n < 0) return CA_ERR_INVALID_ARG;
  if (n == 0) return CA_OK;
  ca::finalize_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      acc, heat, n, inv_count, round_bf16);
  CA_LAUNCH_CHECK();
  return CA_OK;
}

extern "C" ca_status ca_gaussian_blur3(const float* in, float* out, int M, int h, int w,
                                       void* stream) {
  if (!in || !out || in == out || M < 0 || h < 2 || w < 2) return CA_ERR_INVALID_ARG;
  const int64_t n = (int64_t)M * h * w;
  if (n == 0) return CA_OK;
  ca::blur3_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(in, out, M, h, w);
  CA_LAUNCH_CHECK();
  return CA_OK;
}

extern "C" ca_status ca_argmax_mask(const float* heat, int64_t* mask, int B, int C, int P,
                                    int num_background, void* stream) {
  if (!heat || !mask || B < 0 || C <= 0 || P < 0) return CA_ERR_INVALID_ARG;
  const int64_t n = (int64_t)B * P;
  if (n == 0) return CA_OK;
  ca::argmax_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      heat, mask, B, C, P, num_background);
  CA_LAUNCH_CHECK();
  return CA_OK;
}

extern "C" ca_status ca_binary_masks(const float* heat, uint8_t* mask, int M, int P, void* stream) {
  if (!heat || !mask || M < 0 || P <= 0) return CA_ERR_INVALID_ARG;
  if (M == 0) return CA_OK;
  ca::binary_mask_kernel<<<M, 256, 0, (cudaStream_t)stream>>>(heat, mask, P);
  CA_LAUNCH_CHECK();
  return CA_OK;
}
