// Micro-benchmark: MUFU.EX2 issue rate for f32, f16x2 and bf16x2 operands on one SM.
// 16 warps x 4096 iterations x 8 independent chains; elements per cycle per SM.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int MODE>
__global__ void __launch_bounds__(512, 1) k(int iters, long long* out, uint32_t* sink, uint32_t seed) {
  uint32_t v[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) v[j] = seed + threadIdx.x * 8 + j;
  __syncthreads();
  const long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (MODE == 0) { float f = __uint_as_float(v[j]); asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(f)); v[j] = __float_as_uint(f); }
      if (MODE == 1) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(v[j]));
      if (MODE == 2) asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(v[j]));
    }
  }
  const long long t1 = clock64();
  __syncthreads();
  if (threadIdx.x == 0) out[0] = t1 - t0;
  uint32_t a = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) a ^= v[j];
  if (a == 0x1234567u) sink[threadIdx.x] = a;
}

int main() {
  long long* d; uint32_t* sink;
  cudaMalloc(&d, 8); cudaMalloc(&sink, 4096);
  const int iters = 4096;
  const char* names[3] = {"ex2.approx.ftz.f32", "ex2.approx.f16x2", "ex2.approx.ftz.bf16x2"};
  for (int mode = 0; mode < 3; ++mode) {
    if (mode == 0) k<0><<<1, 512>>>(iters, d, sink, 0x3c003c00u);
    if (mode == 1) k<1><<<1, 512>>>(iters, d, sink, 0x3c003c00u);
    if (mode == 2) k<2><<<1, 512>>>(iters, d, sink, 0x3c003c00u);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
    long long c; cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost);
    const double instr = 512.0 * iters * 8;
    printf("%-24s cycles=%lld  thread-instr/cycle/SM=%.2f  elements/cycle/SM=%.2f\n", names[mode], c,
           instr / c, instr * (mode ? 2 : 1) / c);
  }
  return 0;
}
