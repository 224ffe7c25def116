// K3+K4, TMA/tcgen05 variant: heatmap logits as a skinny GEMM streamed at HBM rate.
//
// logits[p, c] = sum_k img[p, k] * con[c, k]  (concept_attention_pipeline.py:57-61) has
// AI = C FLOP/B: HBM-bound, but for C > ~10 the fp32 CUDA cores can no longer keep up with
// 6.5 TB/s, so the contraction is issued as tcgen05.mma (M = 128 image rows, N = C padded to
// 16/32, K chunks of 64 channels) purely to keep the kernel on the memory roofline:
//   warp 0      TMA producer: [128 x 64] image chunk + [NPAD x 64] concept chunk per stage,
//               SWIZZLE_128B, zero fill past P / C, 5-stage mbarrier ring (100 KB in flight)
//   warp 1      MMA issuer: 4 x tcgen05.mma (K = 16) per chunk into NPAD TMEM columns
//   warps 2-5   epilogue: thread == image row: tcgen05.ld the C logits, cluster reduce,
//               softmax / sparsemax / entmax15 in registers, acc[b,c,p] += (coalesced over p)
// Small problems (P = 4096 -> 32 row tiles) are split along K over a thread-block cluster of
// S CTAs; partial logits are reduced through distributed shared memory in rank order, so the
// result is deterministic (no atomics).
#include "ca_common.cuh"
#include "heatmap_norm.cuh"
#include "sm100_ptx.cuh"
#include "tma_host.cuh"

namespace ca {

using namespace ptx;

constexpr int kTcStages = 5;
constexpr int kTcRows = 128;
constexpr int kTcABytes = kTcRows * 128;      // 128 rows x 64 bf16

template <int NPAD>
struct SmemTc {
  union {
    struct {
      uint8_t a[kTcStages][kTcABytes];
      uint8_t b[kTcStages][NPAD * 128];
    } pipe;
    float red[(NPAD == 32 ? 3 : 7)][NPAD][kTcRows];   // peers' partial logits (leader only)
  };
  uint64_t full[kTcStages], empty[kTcStages], acc_done;
  uint32_t tmem_base;
};

struct HmTcParams {
  int B, P, C, seg_len, chunks_per_cta, norm, round_bf16;
  float* acc;
  float* pair; int64_t pair_stride; const int* pair_slot; int pair_slot_off;   // see HmArgs
};

template <int NPAD, int CT, bool kLogitsOnly>
__global__ void __launch_bounds__(192)
heatmap_tc_kernel(const __grid_constant__ CUtensorMap tma0, const __grid_constant__ CUtensorMap tmb0,
                  const __grid_constant__ CUtensorMap tma1, const __grid_constant__ CUtensorMap tmb1,
                  const HmTcParams p0, const HmTcParams p1) {
  // blockIdx.z selects one of two independent problems of the same (B, P, C, K): the output-space
  // and the cross-attention-space maps of one layer go out in ONE launch (gridDim.z = 2)
  const bool second = blockIdx.z != 0;
  const CUtensorMap* tma = second ? &tma1 : &tma0;
  const CUtensorMap* tmb = second ? &tmb1 : &tmb0;
  const HmTcParams& p = second ? p1 : p0;
  extern __shared__ uint8_t smem_raw[];
  SmemTc<NPAD>& sm = *reinterpret_cast<SmemTc<NPAD>*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const uint32_t S = cluster_nctarank(), rank = cluster_ctarank();
  const int tile = blockIdx.x / S, b = blockIdx.y;
  const int row0 = tile * kTcRows;
  const int chunk0 = rank * p.chunks_per_cta, n = p.chunks_per_cta;
  // Every row tile needs the same concept chunks: start each CTA at a different point of its
  // (cyclic) K range so the CTAs of a wave do not hammer the same L2 lines in lockstep.
  const int kshift = (tile * 5 + b * 3) % n;
  constexpr uint32_t kTmemCols = NPAD < 32 ? 32 : NPAD;

  if (threadIdx.x == 0) {
    prefetch_tmap(tma); prefetch_tmap(tmb);
    for (int i = 0; i < kTcStages; ++i) { mbar_init(&sm.full[i], 1); mbar_init(&sm.empty[i], 1); }
    mbar_init(&sm.acc_done, 1);
    fence_barrier_init();
  }
  if (warp == 1) { tmem_alloc(&sm.tmem_base, kTmemCols); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  float z[NPAD];
#pragma unroll
  for (int c = 0; c < NPAD; ++c) z[c] = 0.f;
  if (warp == 0) {
    if (elect_one()) {
      for (int i = 0; i < n; ++i) {
        const int st = i % kTcStages;
        mbar_wait(&sm.empty[st], ((i / kTcStages) & 1) ^ 1);
        int ci = i + kshift; ci = ci >= n ? ci - n : ci;
        const int e = (chunk0 + ci) * 64, seg = e / p.seg_len, off = e - seg * p.seg_len;
        mbar_expect_tx(&sm.full[st], kTcABytes + NPAD * 128);
        tma_load_4d(sm.pipe.a[st], tma, &sm.full[st], off, row0, seg, b);
        tma_load_4d(sm.pipe.b[st], tmb, &sm.full[st], off, 0, seg, b);
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(kTcRows, NPAD, false, false);
      for (int i = 0; i < n; ++i) {
        const int st = i % kTcStages;
        mbar_wait(&sm.full[st], (i / kTcStages) & 1);
        tc_fence_after();
        const uint32_t aa = smem_u32(sm.pipe.a[st]), ba = smem_u32(sm.pipe.b[st]);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          mma_ss(tmem, make_smem_desc(aa + kk * 32, 16, 1024), make_smem_desc(ba + kk * 32, 16, 1024),
                 idesc, i > 0 || kk > 0);
        mma_commit(&sm.empty[st]);
      }
      mma_commit(&sm.acc_done);
    }
  } else {
    mbar_wait(&sm.acc_done, 0);
    tc_fence_after();
    uint32_t r[NPAD];
    const uint32_t taddr = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    if (NPAD == 16) tmem_ld16(taddr, r); else tmem_ld32(taddr, r);
    tmem_wait_ld();
#pragma unroll
    for (int c = 0; c < NPAD; ++c) z[c] = __uint_as_float(r[c]);
  }

  const int row = (warp & 3) * 32 + lane;          // epilogue warps only
  if (S > 1) {
    cluster_arrive(); cluster_wait();              // every CTA of the cluster is past its main loop
    if (warp >= 2 && rank != 0) {
      const uint32_t base = mapa(smem_u32(&sm.red[rank - 1][0][row]), 0);
#pragma unroll
      for (int c = 0; c < NPAD; ++c)
        if (c < p.C) st_cluster_f32(base + c * kTcRows * 4, z[c]);
    }
    cluster_arrive(); cluster_wait();              // partials have landed in the leader
  }
  if (warp >= 2 && rank == 0) {
    for (uint32_t r = 1; r < S; ++r) {
#pragma unroll
      for (int c = 0; c < NPAD; ++c)
        if (c < p.C) z[c] += sm.red[r - 1][c][row];
    }
    float zz[CT];
#pragma unroll
    for (int c = 0; c < CT; ++c) zz[c] = z[c];
    if (!kLogitsOnly) normalise<CT>(zz, p.C, p.norm, p.round_bf16 != 0);
    const int pi = row0 + row;
    if (pi < p.P) {
      float* dst = p.acc + (int64_t)b * p.C * p.P + pi;
      if (kLogitsOnly) {
#pragma unroll
        for (int c = 0; c < CT; ++c) if (c < p.C) dst[(int64_t)c * p.P] = zz[c];
      } else {
        // issue every accumulator load before the first store: the compiler cannot prove the
        // strided stores do not alias the next load and would serialise C DRAM round trips
        float old[CT];
#pragma unroll
        for (int c = 0; c < CT; ++c) old[c] = c < p.C ? __ldcg(dst + (int64_t)c * p.P) : 0.f;
#pragma unroll
        for (int c = 0; c < CT; ++c) if (c < p.C) dst[(int64_t)c * p.P] = old[c] + zz[c];
        if (p.pair) {
          const int slot = (p.pair_slot ? __ldg(p.pair_slot) : 0) + p.pair_slot_off;
          float* pd = p.pair + (int64_t)slot * p.pair_stride + (int64_t)b * p.C * p.P + pi;
#pragma unroll
          for (int c = 0; c < CT; ++c) if (c < p.C) pd[(int64_t)c * p.P] = zz[c];
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, kTmemCols);
}

static bool make_tmap4(CUtensorMap* m, const void* base, uint64_t d0, uint64_t d1, uint64_t d2,
                       uint64_t d3, int64_t s1, int64_t s2, int64_t s3, uint32_t box_rows) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) return false;
  // strides of size-1 dimensions are never used for addressing; give them a legal value
  if (d2 == 1) s2 = s1 * (int64_t)d1;
  if (d3 == 1) s3 = s2 * (int64_t)d2;
  const cuuint64_t dims[4] = {d0, d1, d2, d3};
  const cuuint64_t strides[3] = {(cuuint64_t)s1 * 2, (cuuint64_t)s2 * 2, (cuuint64_t)s3 * 2};
  const cuuint32_t box[4] = {64, box_rows, 1, 1};
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  return enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
             CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template <int NPAD, int CT, bool kLogitsOnly>
static ca_status launch_tc(const CUtensorMap& tma, const CUtensorMap& tmb, const HmTcParams& p,
                           int tiles, int S, cudaStream_t st, const CUtensorMap* tma1 = nullptr,
                           const CUtensorMap* tmb1 = nullptr, const HmTcParams* p1 = nullptr) {
  auto kern = heatmap_tc_kernel<NPAD, CT, kLogitsOnly>;
  const int smem = (int)sizeof(SmemTc<NPAD>) + 1024;
  CA_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(tiles * S, p.B, p1 ? 2 : 1);
  cfg.blockDim = dim3(192);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = S; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  CA_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, tma, tmb, tma1 ? *tma1 : tma, tmb1 ? *tmb1 : tmb, p, p1 ? *p1 : p));
  return CA_OK;
}


ca_status launch_heatmap_tc(const HmArgs& a, bool logits_only, cudaStream_t st) {
  if (a.seg_len % 64 || a.C > 32 || a.P < 1 || a.B > 65535) return CA_ERR_UNSUPPORTED;
  const int chunks = a.nseg * (a.seg_len / 64);
  const int tiles = (a.P + kTcRows - 1) / kTcRows;
  const int npad = a.C <= 16 ? 16 : 32;
  CUtensorMap tma, tmb;
  if (!make_tmap4(&tma, a.img, a.seg_len, a.P, a.nseg, a.B, a.isp, a.iss, a.isb, kTcRows) ||
      !make_tmap4(&tmb, a.con, a.seg_len, a.C, a.nseg, a.B, a.csc, a.css, a.csb, npad))
    return CA_ERR_UNSUPPORTED;
  // split K over a cluster only until the grid covers most of the chip (measured: extra
  // splitting costs more in cluster barriers than it gains once >= 128 CTAs stream)
  int S = 1;
  const int smax = npad == 32 ? 4 : 8;
  while (S < smax && (int64_t)tiles * a.B * S < 128 && chunks % (2 * S) == 0 && chunks / (2 * S) >= 2)
    S *= 2;
  HmTcParams p{a.B, a.P, a.C, a.seg_len, chunks / S, a.norm, a.round_bf16, a.acc,
               a.pair, a.pair_stride, a.pair_slot, a.pair_slot_off};
#define CA_TC(NP, CT)                                                                 \
  (logits_only ? launch_tc<NP, CT, true>(tma, tmb, p, tiles, S, st)                   \
               : launch_tc<NP, CT, false>(tma, tmb, p, tiles, S, st))
  if (a.C <= 4) return CA_TC(16, 4);
  if (a.C <= 8) return CA_TC(16, 8);
  if (a.C <= 16) return CA_TC(16, 16);
  return CA_TC(32, 32);
#undef CA_TC
}

// Two problems of the same (B, P, C, total K, norm) in one launch; CA_ERR_UNSUPPORTED when either
// does not fit the TMA/tcgen05 kernel (the caller then launches them one by one).
ca_status launch_heatmap_tc_dual(const HmArgs& a, const HmArgs& b, cudaStream_t st) {
  if (a.B != b.B || a.P != b.P || a.C != b.C || a.norm != b.norm || a.round_bf16 != b.round_bf16 ||
      a.nseg * a.seg_len != b.nseg * b.seg_len)
    return CA_ERR_UNSUPPORTED;
  if (a.seg_len % 64 || b.seg_len % 64 || a.C > 32 || a.P < 1 || a.B > 65535) return CA_ERR_UNSUPPORTED;
  const int chunks = a.nseg * (a.seg_len / 64);
  const int tiles = (a.P + kTcRows - 1) / kTcRows;
  const int npad = a.C <= 16 ? 16 : 32;
  CUtensorMap ta, tb, ua, ub;
  if (!make_tmap4(&ta, a.img, a.seg_len, a.P, a.nseg, a.B, a.isp, a.iss, a.isb, kTcRows) ||
      !make_tmap4(&tb, a.con, a.seg_len, a.C, a.nseg, a.B, a.csc, a.css, a.csb, npad) ||
      !make_tmap4(&ua, b.img, b.seg_len, b.P, b.nseg, b.B, b.isp, b.iss, b.isb, kTcRows) ||
      !make_tmap4(&ub, b.con, b.seg_len, b.C, b.nseg, b.B, b.csc, b.css, b.csb, npad))
    return CA_ERR_UNSUPPORTED;
  int S = 1;                                           // twice the CTAs of a single launch: split K less
  const int smax = npad == 32 ? 4 : 8;
  while (S < smax && (int64_t)tiles * a.B * S * 2 < 128 && chunks % (2 * S) == 0 && chunks / (2 * S) >= 2)
    S *= 2;
  HmTcParams pa{a.B, a.P, a.C, a.seg_len, chunks / S, a.norm, a.round_bf16, a.acc,
                a.pair, a.pair_stride, a.pair_slot, a.pair_slot_off};
  HmTcParams pb{b.B, b.P, b.C, b.seg_len, chunks / S, b.norm, b.round_bf16, b.acc,
                b.pair, b.pair_stride, b.pair_slot, b.pair_slot_off};
#define CA_TC2(NP, CT) launch_tc<NP, CT, false>(ta, tb, pa, tiles, S, st, &ua, &ub, &pb)
  if (a.C <= 4) return CA_TC2(16, 4);
  if (a.C <= 8) return CA_TC2(16, 8);
  if (a.C <= 16) return CA_TC2(16, 16);
  return CA_TC2(32, 32);
#undef CA_TC2
}

}  // namespace ca

