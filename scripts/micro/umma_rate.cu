// tcgen05.mma issue-rate and TMEM-read microbenchmark behind the K2 design decisions
// (profiles/r02_umma_rate.md).  Question 1: how many cycles does one K=16 bf16 tcgen05.mma take at
// M=128 (cta_group::1) or M=256 (cta_group::2) when both operands come from shared memory (SS) or
// A comes from TMEM (TS), at N = 64 / 128 / 256?  One CTA (pair) per SM on all SMs; one thread issues
// groups of 8 MMAs back to back, commits once, waits.  Question 2: how fast can 1 / 2 / 4 / 8 warps
// pull a 128x128 fp32 tile out of TMEM (tcgen05.ld.32x32b.x32)?
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o scripts/micro/umma_rate scripts/micro/umma_rate.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <cuda.h>
#include "../../conceptattention_b200/csrc/sm100_ptx.cuh"

namespace {
using namespace ca::ptx;
constexpr int kA = 0, kB = 65536, kBar = 196608;   // A 64 KB, B 128 KB (N up to 256), barrier
constexpr int kSmem = kBar + 64;

__device__ __forceinline__ void mma_ss2(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, bool acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
               ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"((uint32_t)acc) : "memory");
}
__device__ __forceinline__ void mma_ts2(uint32_t d, uint32_t a, uint64_t b, uint32_t idesc, bool acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
               ::"r"(d), "r"(a), "l"(b), "r"(idesc), "r"((uint32_t)acc) : "memory");
}
__device__ __forceinline__ void commit2(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}

// kMode: 0 SS, 1 TS, 2 alternate groups SS / TS (the attention pattern: QK from smem, PV with P in TMEM)
template <int kCtas, int kMode, int N>
__global__ void __launch_bounds__(128, 1) rate_kernel(int groups, long long* cycles) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint32_t tmem_base;
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + kBar);
  const int warp = threadIdx.x >> 5;
  uint32_t rank = 0;
  if (kCtas == 2) rank = cluster_ctarank();
  for (int i = threadIdx.x; i < kBar / 4; i += 128) {
    uint32_t h = (i * 2654435761u) ^ (blockIdx.x * 40503u);
    reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u | (h & 0x007f007fu);   // bf16 pairs near 0.0078
  }
  if (threadIdx.x == 0) { mbar_init(bar, 1); fence_barrier_init(); }
  if (warp == 0) {
    if (kCtas == 2)
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base)), "r"(512) : "memory");
    else
      tmem_alloc(&tmem_base, 512);
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  if (kCtas == 2) { cluster_arrive(); cluster_wait(); }
  tc_fence_after();
  const uint32_t tmem = tmem_base;
  {  // A operand for the TS forms: columns 384..447 (128 rows x 128 bf16)
    uint32_t r[32];
    for (int i = 0; i < 32; ++i) r[i] = 0x3c003c00u | ((threadIdx.x * 131 + i * 7) & 0x7f);
    tmem_st32(tmem + ((uint32_t)(warp * 32) << 16) + 384, r);
    tmem_st32(tmem + ((uint32_t)(warp * 32) << 16) + 416, r);
    tmem_wait_st();
  }
  tc_fence_before();
  __syncthreads();
  if (kCtas == 2) { cluster_arrive(); cluster_wait(); }
  tc_fence_after();
  long long t0 = 0, t1 = 0;
  if (threadIdx.x == 0 && rank == 0) {
    const uint32_t idesc_k = make_idesc_bf16(128 * kCtas, N, false, false);   // B K-major (like K)
    const uint32_t idesc_mn = make_idesc_bf16(128 * kCtas, N, false, true);   // B MN-major (like V)
    const uint64_t da = make_smem_desc(smem_u32(smem + kA), 16, 1024);
    const uint64_t dbk = make_smem_desc(smem_u32(smem + kB), 16, 1024);
    const uint64_t dbv = make_smem_desc(smem_u32(smem + kB), 16384, 1024);
    t0 = clock64();
#pragma unroll 1
    for (int g = 0; g < groups; ++g) {
      const uint32_t d = tmem + (g & 1) * (N > 128 ? 0 : 128);
      if (kMode == 1 || (kMode == 2 && (g & 2))) {
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          if (kCtas == 2) mma_ts2(d, tmem + 384 + kk * 8, dbv + ((kk * 2048) >> 4), idesc_mn, kk > 0);
          else mma_ts(d, tmem + 384 + kk * 8, dbv + ((kk * 2048) >> 4), idesc_mn, kk > 0);
        }
      } else {
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          const uint32_t off = ((kk >> 2) * 16384 + (kk & 3) * 32) >> 4;
          if (kCtas == 2) mma_ss2(d, da + off, dbk + off, idesc_k, kk > 0);
          else mma_ss(d, da + off, dbk + off, idesc_k, kk > 0);
        }
      }
    }
    if (kCtas == 2) commit2(bar); else mma_commit(bar);
  }
  if (threadIdx.x == 0) {
    mbar_wait(bar, 0);
    t1 = clock64();
    if (rank == 0) cycles[blockIdx.x / kCtas] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (kCtas == 2) { cluster_arrive(); cluster_wait(); }
  if (warp == 0) {
    if (kCtas == 2)
      asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
    else
      tmem_dealloc(tmem, 512);
  }
}

// `nwarps` warps (warp w reads lanes 32 (w % 4) ..) each issue `iters` x 4 tcgen05.ld.32x32b.x32
__global__ void __launch_bounds__(256, 1) ldtm_kernel(int nwarps, int iters, long long* cycles, float* sink) {
  __shared__ uint32_t tmem_base;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc(&tmem_base, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
  float acc = 0.f;
  long long t0 = clock64();
  if (warp < nwarps) {
    for (int it = 0; it < iters; ++it) {
      uint32_t r[128];
      tmem_ld32(tmem + 0, r); tmem_ld32(tmem + 32, r + 32); tmem_ld32(tmem + 64, r + 64); tmem_ld32(tmem + 96, r + 96);
      tmem_wait_ld32(r); tmem_wait_ld32(r + 32); tmem_wait_ld32(r + 64); tmem_wait_ld32(r + 96);
#pragma unroll
      for (int i = 0; i < 128; i += 16) acc += __uint_as_float(r[i]);
    }
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  if (acc == 123.456f) sink[0] = acc;
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, 512);
}

template <int kCtas, int kMode, int N> float run(int groups, int sms, long long* dcyc) {
  auto kern = rate_kernel<kCtas, kMode, N>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(sms / kCtas * kCtas); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = kSmem;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = kCtas; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f;
  for (int it = 0; it < 4; ++it) {
    cudaEventRecord(e0);
    cudaError_t err = cudaLaunchKernelEx(&cfg, kern, groups, dcyc);
    cudaEventRecord(e1);
    if (err != cudaSuccess || cudaEventSynchronize(e1) != cudaSuccess) {
      printf("launch failed: %s\n", cudaGetErrorString(cudaGetLastError())); exit(1);
    }
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (it > 0 && ms < best) best = ms;
  }
  return best;
}
}  // namespace

int main() {
  cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
  const int sms = prop.multiProcessorCount, groups = 8192;
  long long* dcyc; cudaMalloc(&dcyc, sizeof(long long) * sms);
  long long* h = (long long*)malloc(sizeof(long long) * sms);
  const char* names[3] = {"SS", "TS", "SS/TS alternating"};
  printf("| cta_group | form | N | cycles per K=16 MMA (avg over CTAs) | N/2 | smem operand bytes per SM per MMA | TFLOP/s all SMs (8 ms burst) |\n|---|---|---|---|---|---|---|\n");
  auto row = [&](int ctas, int mode, int N, float ms) {
    cudaMemcpy(h, dcyc, sizeof(long long) * (sms / ctas), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < sms / ctas; ++i) avg += h[i]; avg /= (sms / ctas);
    const double per = avg / (groups * 8.0);
    const double flops = 2.0 * 128 * ctas * N * 16 * groups * 8.0 * (sms / ctas);
    const int a_b = 128 * 16 * 2, b_b = N / ctas * 16 * 2;
    const int bytes = mode == 0 ? a_b + b_b : mode == 1 ? b_b : (a_b + 2 * b_b) / 2;
    printf("| %d | %s | %d | %.1f | %d | %d | %.0f |\n", ctas, names[mode], N, per, N / 2, bytes,
           flops / (ms * 1e-3) / 1e12);
  };
#define ROW(C, M, N) row(C, M, N, run<C, M, N>(groups, sms, dcyc))
  ROW(1, 0, 64); ROW(1, 0, 128); ROW(1, 0, 256); ROW(1, 1, 64); ROW(1, 1, 128); ROW(1, 2, 64); ROW(1, 2, 128);
  ROW(2, 0, 64); ROW(2, 0, 128); ROW(2, 0, 256); ROW(2, 1, 64); ROW(2, 1, 128); ROW(2, 2, 64); ROW(2, 2, 128);
  printf("\n| warps reading TMEM | cycles per 4 x tcgen05.ld.32x32b.x32 (16 KB per warp) | bytes per cycle per SM |\n|---|---|---|\n");
  float* sink; cudaMalloc(&sink, 4);
  for (int nw : {1, 2, 4, 8}) {
    const int iters = 4096;
    ldtm_kernel<<<sms, 256>>>(nw, iters, dcyc, sink);
    ldtm_kernel<<<sms, 256>>>(nw, iters, dcyc, sink);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("ldtm failed\n"); return 1; }
    cudaMemcpy(h, dcyc, sizeof(long long) * sms, cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < sms; ++i) avg += h[i]; avg /= sms;
    printf("| %d | %.1f | %.1f |\n", nw, avg / iters, nw * 16384.0 * iters / avg);
  }
  return 0;
}
