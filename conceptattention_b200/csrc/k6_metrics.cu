// K6: segmentation-metric counters on device.  The reference moves predictions
// and labels to the host and counts with numpy (concept_attention/utils.py:71-108):
//   batch_pix_accuracy          utils.py:71-85    pixel_correct / pixel_labeled
//   batch_intersection_union    utils.py:88-108   np.histogram(bins=nclass, range=(1, nclass)) x3
// Both shift labels by +1 so that the ignore label -1 becomes 0 and drops out of every count.
// One pass over the two label maps produces all of them: counts = [correct, labeled,
// inter[nclass], pred[nclass], lab[nclass]] as uint64.  Labels are integer-valued (float storage, as
// the reference passes float tensors); a value v in 1..nclass lands in histogram bin v-1, which is
// what np.histogram's equal-width bins over (1, nclass) do for integers, anything else is dropped.
// Measured (scripts/bench_seqpar.py, uniformly random labels = worst case for contention): 0.32-0.47 of
// the HBM copy bandwidth (64 M pixels, 21 classes: 177 us); the shared-memory atomics bound it.  A
// warp-aggregated form (__match_any_sync + one atomic per distinct bin) was 3x SLOWER on random labels
// (581 us) and is not used.
#include "ca_common.cuh"

namespace ca {

constexpr int kMetricSmemClasses = 1024;

__device__ __forceinline__ int label_bin(float v, int nclass) {
  // integer-valued v; outside 1..nclass (the shifted ignore label 0 included) -> no bin
  if (!(v >= 1.f) || v > (float)nclass) return -1;
  return (int)v - 1;
}

template <bool kSmem>
__global__ void __launch_bounds__(256)
seg_counts_kernel(const float* __restrict__ predict, const float* __restrict__ target, int64_t n,
                  int nclass, unsigned long long* __restrict__ counts) {
  extern __shared__ unsigned int hist[];   // [3][nclass] when kSmem
  __shared__ unsigned int blk[2];
  if (kSmem)
    for (int i = threadIdx.x; i < 3 * nclass; i += blockDim.x) hist[i] = 0u;
  if (threadIdx.x < 2) blk[threadIdx.x] = 0u;
  __syncthreads();

  unsigned long long* g_inter = counts + 2;
  unsigned long long* g_pred = g_inter + nclass;
  unsigned long long* g_lab = g_pred + nclass;
  unsigned int correct = 0, labeled = 0;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float p1 = predict[i] + 1.f, t = target[i] + 1.f;
    const bool lab = t > 0.f;
    const float p = lab ? p1 : 0.f;                 // predict * (target > 0)
    const bool hit = lab && p == t;
    labeled += lab;
    correct += hit;
    const int bp = label_bin(p, nclass), bt = label_bin(t, nclass);
    const int bi = hit ? bp : -1;                   // intersection = predict * (predict == target)
    if (kSmem) {
      if (bp >= 0) atomicAdd(&hist[nclass + bp], 1u);
      if (bt >= 0) atomicAdd(&hist[2 * nclass + bt], 1u);
      if (bi >= 0) atomicAdd(&hist[bi], 1u);
    } else {
      if (bp >= 0) atomicAdd(&g_pred[bp], 1ull);
      if (bt >= 0) atomicAdd(&g_lab[bt], 1ull);
      if (bi >= 0) atomicAdd(&g_inter[bi], 1ull);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    correct += __shfl_xor_sync(0xffffffffu, correct, o);
    labeled += __shfl_xor_sync(0xffffffffu, labeled, o);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(&blk[0], correct);
    atomicAdd(&blk[1], labeled);
  }
  __syncthreads();
  if (threadIdx.x < 2 && blk[threadIdx.x]) atomicAdd(&counts[threadIdx.x], (unsigned long long)blk[threadIdx.x]);
  if (kSmem)
    for (int i = threadIdx.x; i < 3 * nclass; i += blockDim.x)
      if (hist[i]) atomicAdd(&g_inter[i], (unsigned long long)hist[i]);
}

}  // namespace ca

extern "C" ca_status ca_segmentation_counts(const float* predict, const float* target, int64_t n,
                                            int nclass, uint64_t* counts, void* stream) {
  if (!counts || nclass <= 0 || n < 0 || (n > 0 && (!predict || !target))) return CA_ERR_INVALID_ARG;
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t e = cudaMemsetAsync(counts, 0, sizeof(uint64_t) * (2 + 3 * (size_t)nclass), s);
  if (e != cudaSuccess) return CA_ERR_CUDA;
  if (n == 0) return CA_OK;
  // each thread handles at most 2^20 elements of a block's < 2^32 share: u32 block counters are safe
  int64_t blocks = (n + 256 * 16 - 1) / (256 * 16);
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (n / (blocks * 256) >= (1 << 20)) return CA_ERR_INVALID_ARG;
  auto* c = reinterpret_cast<unsigned long long*>(counts);
  if (nclass <= ca::kMetricSmemClasses)
    ca::seg_counts_kernel<true><<<(unsigned)blocks, 256, 3 * nclass * sizeof(unsigned int), s>>>(
        predict, target, n, nclass, c);
  else
    ca::seg_counts_kernel<false><<<(unsigned)blocks, 256, 0, s>>>(predict, target, n, nclass, c);
  CA_LAUNCH_CHECK();
  return CA_OK;
}
