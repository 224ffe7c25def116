// Micro-benchmark: TMEM read bandwidth seen by tcgen05.ld (LDTM) on one SM.
// W warps (warp w reads lanes 32*(w%4)..+31) each issue `iters` x tcgen05.ld.32x32b.x32 (4 KB
// per instruction) over the 512 allocated columns; cycles are counted with clock64 around the
// loop of warp 0 after a CTA barrier.  Prints bytes per cycle per SM for W = 4, 8, 16.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tmem_ld_bw tmem_ld_bw.cu && ./tmem_ld_bw
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(512, 1) k(int warps, int iters, long long* out_cycles, uint32_t* sink) {
  __shared__ uint32_t tmem_base;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t base = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
  uint32_t acc = 0;
  long long t0 = 0, t1 = 0;
  __syncthreads();
  if (warp < warps) {
    t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      uint32_t r[32];
      const uint32_t addr = base + ((i * 32) & 511);
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x32.b32"
          "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15,"
          " %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
          : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
            "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
            "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]),
            "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
            "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
          : "r"(addr) : "memory");
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
      for (int j = 0; j < 32; ++j) acc ^= r[j];
    }
    t1 = clock64();
  }
  __syncthreads();
  if (threadIdx.x == 0) out_cycles[0] = t1 - t0;
  if (acc == 0x12345678u) sink[threadIdx.x] = acc;
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
}

int main() {
  long long* d; uint32_t* sink;
  cudaMalloc(&d, 8); cudaMalloc(&sink, 4096);
  const int iters = 4096;
  for (int warps : {1, 4, 8, 16}) {
    k<<<1, 512>>>(warps, iters, d, sink);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
    long long c; cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost);
    const double bytes = (double)warps * iters * 4096.0;
    printf("warps=%2d  cycles=%lld  cycles/instr(warp0)=%.1f  SM-wide bytes/cycle=%.1f\n", warps, c,
           (double)c / iters, bytes / (double)c);
  }
  return 0;
}
