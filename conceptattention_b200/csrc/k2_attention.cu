// K2: one-way joint MM-DiT attention on tcgen05 / TMEM / TMA (sm_100a).
//
// Replaces concept_attention/modified_double_stream_block.py:106-176 — two
// F.scaled_dot_product_attention calls over cat(txt,img) and cat(concept,img) —
// with ONE launch over a packed sequence [img | txt | concept] per (batch, head):
//   query rows [0, P+T)       see keys [0, P+T)
//   query rows [P+T, P+T+C)   see keys [0, P) U [P+T, P+T+C)   (ablation modes trim this)
// so the image stream is bit-identical to attention without concepts and the
// (C+P)^2 - C(C+P) image rows of the reference's second call are never computed.
//
// Tiling: 128 query rows x 128 keys x d=128, bf16 operands, fp32 accumulation.
//   S = Q K^T : tcgen05.mma kind::f16, A = Q tile (smem, K-major SW128),
//               B = K tile (smem, K-major SW128), D = S in TMEM (128 lanes x 128 cols)
//   O += P V  : A = P (bf16), B = V tile (smem, MN-major SW128), D = O in TMEM
// Tiles arrive by TMA (cp.async.bulk.tensor, 128B swizzle, zero OOB fill) and are
// tracked with mbarriers; MMA completion is signalled with tcgen05.commit.
#include "ca_common.cuh"
#include "sm100_ptx.cuh"
#include "tma_host.cuh"

namespace ca {

using namespace ptx;

constexpr int kD = CA_HEAD_DIM;        // 128
constexpr int kBM = 128;               // query rows per tile
constexpr int kBN = 128;               // keys per tile
constexpr int kTileBytes = kBM * kD * 2;          // 32 KB: one 128x128 bf16 tile
constexpr int kHalfBytes = kTileBytes / 2;        // 16 KB: one 128x64 swizzle block
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kRescaleThreshold = 8.0f;         // lazy rescale: keep the stale max while
                                                  // exp2 arguments stay <= 8 (P <= 256)

struct AttnParams {
  int B, H, P, T, C, mode;
  int n_main_tiles;                    // ceil((P+T)/128)
  __nv_bfloat16* out_img;              // [B,P,H*128]
  __nv_bfloat16* out_txtc;             // [B,T+C,H*128]
  float scale_log2;                    // log2(e)/sqrt(128)
  // flat 1-D grid: [ concept units | whole main units | stream-K tail pieces ] — the concept units
  // (one 128-row tile over all P+C keys: long and not splittable here) start first, the short tail
  // pieces fill the SMs last
  int n_pairs;                         // main units (query-tile pairs) per (b,h)
  int n_dp;                            // main units run whole, one CTA each
  int sk_units;                        // tail main units whose key steps are spread over sk_sms ranges
  int sk_sms;                          // number of ranges (2 CTAs per range, possibly empty); 0 = no split
  int n_concept;                       // concept units (one per (b,h)) or 0
  float* ws_part;                      // [2*sk_sms] partial slots (kPartFloats each)
  int* ws_count;                       // [sk_units] arrival counters (zeroed by the launcher)
};

// one partial slot: per query tile an unnormalised fp32 O (stored [col/4][row][4] so that the 32
// rows of a warp form 512-byte runs), the row max it is scaled to and the row sum
constexpr int kPartTile = kBM * kD + 2 * kBM;
constexpr int kPartFloats = 2 * kPartTile;

// Key ranges visible to a query tile: up to two half-open row segments.
struct KeySegs { int s0, e0, s1, e1; };

__device__ __forceinline__ KeySegs key_segments(const AttnParams& p, bool concept_tile) {
  const int PT = p.P + p.T;
  if (!concept_tile) return {0, PT, 0, 0};
  if (p.mode == CA_ATTN_CROSS_ONLY) return {0, p.P, 0, 0};
  if (p.mode == CA_ATTN_SELF_ONLY) return {PT, PT + p.C, 0, 0};
  return {0, p.P, PT, PT + p.C};
}

__device__ __forceinline__ int tiles_of(int s, int e) { return (e - s + kBN - 1) / kBN; }

// key tile j of the segments -> first row and number of valid rows
__device__ __forceinline__ void key_tile(const KeySegs& ks, int n0, int j, int& row0, int& valid) {
  if (j < n0) { row0 = ks.s0 + j * kBN; valid = min(kBN, ks.e0 - row0); }
  else        { row0 = ks.s1 + (j - n0) * kBN; valid = min(kBN, ks.e1 - row0); }
}

// Write 8 consecutive bf16 (one 16-byte chunk) of row r, columns [c8*8, c8*8+8) into a
// K-major SWIZZLE_128B operand tile made of two 128x64 blocks.
__device__ __forceinline__ void st_swizzled_chunk(uint8_t* tile, int r, int c8, uint4 val) {
  const int blk = c8 >> 3, cc = c8 & 7;
  *reinterpret_cast<uint4*>(tile + blk * kHalfBytes + r * 128 + ((cc ^ (r & 7)) << 4)) = val;
}

__device__ __forceinline__ void store_out_row(const AttnParams& p, int b, int h, int row,
                                              const float* o, float inv_l, int c0, int ncols) {
  // row is an index into the packed sequence; img rows -> out_img, the rest -> out_txtc
  const int HD = p.H * kD;
  __nv_bfloat16* dst = row < p.P
      ? p.out_img + ((int64_t)b * p.P + row) * HD + h * kD
      : p.out_txtc + ((int64_t)b * (p.T + p.C) + (row - p.P)) * HD + h * kD;
#pragma unroll
  for (int c = 0; c < ncols; c += 8) {
    float t[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) t[j] = o[c + j] * inv_l;
    *reinterpret_cast<uint4*>(dst + c0 + c) = pack8(t);
  }
}

// =============================================================================
// v0: one 128-row query tile per CTA, 128 threads, synchronous softmax.  Thread 0
// issues TMA and MMA; every thread owns one query row (TMEM lane == thread).
// Kept as the simple, obviously-correct implementation the pipelined kernel is
// validated against.
// =============================================================================
struct SmemV0 {
  uint8_t q[kTileBytes];
  uint8_t k[2][kTileBytes];
  uint8_t v[2][kTileBytes];
  uint8_t p[kTileBytes];
  uint64_t bar_q, bar_k[2], bar_v[2], bar_s, bar_o;
  uint32_t tmem_base;
};

__global__ void __launch_bounds__(128, 1)
attn_v0_kernel(const __grid_constant__ CUtensorMap tmq, const __grid_constant__ CUtensorMap tmk,
               const __grid_constant__ CUtensorMap tmv, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  SmemV0& sm = *reinterpret_cast<SmemV0*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int tid = threadIdx.x, warp = tid >> 5;
  const int bh = blockIdx.y, b = bh / p.H, h = bh - b * p.H;
  const bool concept_tile = (int)blockIdx.x >= p.n_main_tiles;
  const int PT = p.P + p.T;
  const int q_row0 = concept_tile ? PT : blockIdx.x * kBM;
  const int q_rows = concept_tile ? p.C : min(kBM, PT - q_row0);
  const KeySegs ks = key_segments(p, concept_tile);
  const int n0 = tiles_of(ks.s0, ks.e0), n_tiles = n0 + tiles_of(ks.s1, ks.e1);

  if (tid == 0) {
    prefetch_tmap(&tmq); prefetch_tmap(&tmk); prefetch_tmap(&tmv);
    mbar_init(&sm.bar_q, 1);
    for (int i = 0; i < 2; ++i) { mbar_init(&sm.bar_k[i], 1); mbar_init(&sm.bar_v[i], 1); }
    mbar_init(&sm.bar_s, 1); mbar_init(&sm.bar_o, 1);
    fence_barrier_init();
  }
  if (warp == 0) { tmem_alloc(&sm.tmem_base, 256); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;
  const uint32_t t_s = tmem, t_o = tmem + 128;
  const uint32_t lane_off = (uint32_t)(warp * 32) << 16;

  auto load_k = [&](int j) {
    int row0, valid; key_tile(ks, n0, j, row0, valid);
    mbar_expect_tx(&sm.bar_k[j & 1], kTileBytes);
    tma_load_3d(sm.k[j & 1], &tmk, &sm.bar_k[j & 1], 0, row0, bh);
    tma_load_3d(sm.k[j & 1] + kHalfBytes, &tmk, &sm.bar_k[j & 1], 64, row0, bh);
  };
  auto load_v = [&](int j) {
    int row0, valid; key_tile(ks, n0, j, row0, valid);
    mbar_expect_tx(&sm.bar_v[j & 1], kTileBytes);
    tma_load_3d(sm.v[j & 1], &tmv, &sm.bar_v[j & 1], 0, row0, bh);
    tma_load_3d(sm.v[j & 1] + kHalfBytes, &tmv, &sm.bar_v[j & 1], 64, row0, bh);
  };
  if (tid == 0) {
    mbar_expect_tx(&sm.bar_q, kTileBytes);
    tma_load_3d(sm.q, &tmq, &sm.bar_q, 0, q_row0, bh);
    tma_load_3d(sm.q + kHalfBytes, &tmq, &sm.bar_q, 64, q_row0, bh);
    load_k(0); load_v(0);
  }

  constexpr uint32_t idesc_qk = make_idesc_bf16(kBM, kBN, false, false);
  constexpr uint32_t idesc_pv = make_idesc_bf16(kBM, kD, false, true);
  float m_used = -INFINITY, l_sum = 0.f;

  for (int j = 0; j < n_tiles; ++j) {
    const int st = j & 1;
    int row0, valid; key_tile(ks, n0, j, row0, valid);
    if (tid == 0) {
      if (j == 0) mbar_wait(&sm.bar_q, 0);
      mbar_wait(&sm.bar_k[st], (j >> 1) & 1);
      tc_fence_after();
#pragma unroll
      for (int kk = 0; kk < kD / 16; ++kk) {
        const uint32_t off = (kk >> 2) * kHalfBytes + (kk & 3) * 32;
        mma_ss(t_s, make_smem_desc(smem_u32(sm.q) + off, 16, 1024),
               make_smem_desc(smem_u32(sm.k[st]) + off, 16, 1024), idesc_qk, kk > 0);
      }
      mma_commit(&sm.bar_s);
    }
    mbar_wait(&sm.bar_s, j & 1);
    tc_fence_after();
    if (tid == 0 && j + 1 < n_tiles) load_k(j + 1);     // K stage st^1 was released by QK(j-1)

    // ---- softmax of this thread's row -------------------------------------------
    float s[kBN];
#pragma unroll
    for (int c = 0; c < kBN; c += 32) {
      uint32_t r[32];
      tmem_ld32(t_s + lane_off + c, r);
      tmem_wait_ld();
#pragma unroll
      for (int i = 0; i < 32; ++i) s[c + i] = __uint_as_float(r[i]);
    }
    float mx = -INFINITY;
#pragma unroll
    for (int c = 0; c < kBN; ++c) { if (c >= valid) s[c] = -INFINITY; mx = fmaxf(mx, s[c]); }
    const bool grow = (mx - m_used) * p.scale_log2 > kRescaleThreshold;   // also true when m_used = -inf
    const float m_new = grow ? mx : m_used;
    const float alpha = grow ? ex2((m_used - m_new) * p.scale_log2) : 1.f;
    const bool warp_rescale = __any_sync(0xffffffffu, grow);

    if (j > 0) {                                   // PV(j-1) done: O and the P tile are free
      mbar_wait(&sm.bar_o, (j - 1) & 1);
      tc_fence_after();
      if (tid == 0 && j + 1 < n_tiles) load_v(j + 1);
      if (warp_rescale) {
#pragma unroll
        for (int c = 0; c < kD; c += 32) {
          uint32_t r[32];
          tmem_ld32(t_o + lane_off + c, r);
          tmem_wait_ld();
#pragma unroll
          for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
          tmem_st32(t_o + lane_off + c, r);
        }
        tmem_wait_st();
      }
    } else if (tid == 0 && n_tiles > 1) {
      load_v(1);
    }
    m_used = m_new;
    l_sum *= alpha;
    const float mb = m_used * p.scale_log2;
    float rs = 0.f;
#pragma unroll
    for (int c8 = 0; c8 < kBN / 8; ++c8) {
      float e[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) { e[i] = ex2(s[c8 * 8 + i] * p.scale_log2 - mb); rs += e[i]; }
      st_swizzled_chunk(sm.p, tid, c8, pack8(e));
    }
    l_sum += rs;
    fence_proxy_async_smem();                       // generic-proxy P writes -> async proxy (MMA)
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      mbar_wait(&sm.bar_v[st], (j >> 1) & 1);
#pragma unroll
      for (int kk = 0; kk < kBN / 16; ++kk) {
        const uint32_t aoff = (kk >> 2) * kHalfBytes + (kk & 3) * 32;
        mma_ss(t_o, make_smem_desc(smem_u32(sm.p) + aoff, 16, 1024),
               make_smem_desc(smem_u32(sm.v[st]) + kk * 2048, kHalfBytes, 1024), idesc_pv,
               j > 0 || kk > 0);
      }
      mma_commit(&sm.bar_o);
    }
  }

  // ---- epilogue: O / l -> bf16 "B L (H D)" -------------------------------------------
  mbar_wait(&sm.bar_o, (n_tiles - 1) & 1);
  tc_fence_after();
  const float inv_l = 1.f / l_sum;
#pragma unroll
  for (int c = 0; c < kD; c += 32) {
    uint32_t r[32];
    tmem_ld32(t_o + lane_off + c, r);
    tmem_wait_ld();
    if (tid < q_rows) {
      float o[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) o[i] = __uint_as_float(r[i]);
      store_out_row(p, b, h, q_row0 + tid, o, inv_l, c, 32);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

// =============================================================================
// v1: warp-specialised kernel.  One CTA = TWO 128-row query tiles of one (b,h)
// (or one tile: odd tail / the concept tile), 384 threads:
//   warp 0      TMA producer: Q tiles once, then K(j), V(j) through a 4-slot ring
//   warp 1      MMA issuer (one elected lane): S_t = Q_t K^T, O_t += P_t V
//   warp 2      TMEM allocator (512 columns: S0 S1 O0 O1), warp 3 idle
//   warps 4-7   softmax warpgroup of tile 0: thread == query row == TMEM lane
//   warps 8-11  softmax warpgroup of tile 1
// P_t (bf16) is written back over the first 64 columns of S_t in TMEM and consumed
// as the A operand of the PV MMA straight from TMEM (no shared-memory round trip).
// The two tiles ping-pong: while warpgroup t runs exp2 on S_t(j), the tensor core
// runs PV_{1-t}(j) and QK_{1-t}(j+1).  O is rescaled lazily (only when a row max
// grows by more than 2^8) by the softmax warpgroup itself, which is safe without an
// extra barrier because the commit that publishes S_t(j) also covers PV_t(j-1).
// =============================================================================
constexpr int kKvStages = 4;

struct SmemV1 {
  uint8_t q[2][kTileBytes];
  uint8_t kv[kKvStages][kTileBytes];
  uint64_t q_full[2], kv_full[kKvStages], kv_empty[kKvStages], s_full[2], p_full[2], o_done;
  uint32_t tmem_base;
  float xchg[2][2][2][kBM];     // v3/v4: [pair][parity][half][row] row max / row sum exchange
};

struct UnitV1 {
  bool concept_tile;
  int n_q;            // query tiles in this CTA (1 or 2)
  int q_row0[2], q_rows[2];
  KeySegs ks;
  int n0, n_tiles;    // key tiles of the unit: n0 in the first segment, n_tiles in total
  int bh;
  int j_begin, n_local;   // key tiles [j_begin, j_begin + n_local) are processed by this CTA
  int part_slot;      // >= 0: write a partial into this slot and take part in the fix-up
  int sk_unit;        // index of the tail unit (fix-up bookkeeping)
};

// tail range i covers key steps [lo, hi) of the concatenated tail units (n steps each)
__host__ __device__ __forceinline__ void sk_range(const AttnParams& p, int n, int i, int64_t& lo, int64_t& hi) {
  const int64_t total = (int64_t)p.sk_units * n;
  lo = (int64_t)i * total / p.sk_sms;
  hi = (int64_t)(i + 1) * total / p.sk_sms;
}
// piece k (0/1) of range i: [lo, hi) clipped to the first / second unit it touches
__host__ __device__ __forceinline__ bool sk_piece(const AttnParams& p, int n, int i, int k, int& unit,
                                         int& j_begin, int& n_local) {
  int64_t lo, hi;
  sk_range(p, n, i, lo, hi);
  if (lo >= hi) return false;
  const int64_t a0 = lo / n, cut = (a0 + 1) * n;
  int64_t s0, s1;
  if (k == 0) { s0 = lo; s1 = hi < cut ? hi : cut; unit = (int)a0; }
  else        { s0 = cut; s1 = hi; unit = (int)a0 + 1; }
  if (s0 >= s1) return false;
  j_begin = (int)(s0 - (int64_t)unit * n);
  n_local = (int)(s1 - s0);
  return true;
}

__device__ __forceinline__ UnitV1 decode_unit(const AttnParams& p, int cta) {
  UnitV1 u;
  const int PT = p.P + p.T;
  const int n_seg = 2 * p.sk_sms;
  u.part_slot = -1; u.sk_unit = -1; u.j_begin = 0;
  (void)n_seg;
  u.concept_tile = cta < p.n_concept;
  int unit = 0;                                           // main unit index over all (b,h)
  bool empty = false;
  if (u.concept_tile) {
    u.bh = cta;
    u.n_q = 1;
    u.q_row0[0] = PT; u.q_rows[0] = p.C; u.q_row0[1] = PT; u.q_rows[1] = 0;
  } else {
    const int n_main = tiles_of(0, PT);
    cta -= p.n_concept;
    if (cta < p.n_dp) {
      unit = cta;
      u.n_local = n_main;
    } else {
      const int s = cta - p.n_dp;
      int a = 0;
      empty = !sk_piece(p, n_main, s >> 1, s & 1, a, u.j_begin, u.n_local);
      unit = p.n_dp + a;
      u.sk_unit = a;
      if (!empty && !(u.j_begin == 0 && u.n_local == n_main)) u.part_slot = s;
    }
    u.bh = unit / p.n_pairs;
    const int pair = unit - u.bh * p.n_pairs;
    u.n_q = min(2, p.n_main_tiles - 2 * pair);
#pragma unroll
    for (int t = 0; t < 2; ++t) {
      u.q_row0[t] = (2 * pair + t) * kBM;
      u.q_rows[t] = max(0, min(kBM, PT - u.q_row0[t]));
    }
  }
  u.ks = key_segments(p, u.concept_tile);
  u.n0 = tiles_of(u.ks.s0, u.ks.e0);
  u.n_tiles = u.n0 + tiles_of(u.ks.s1, u.ks.e1);
  if (u.concept_tile) u.n_local = u.n_tiles;
  if (empty) u.n_local = 0;
  return u;
}

__device__ __forceinline__ void v1_producer(SmemV1& sm, const CUtensorMap* tmq,
                                            const CUtensorMap* tmk, const CUtensorMap* tmv,
                                            const UnitV1& u, int bh) {
  int item = 0;
  auto load_tile = [&](uint8_t* dst, const CUtensorMap* tm, uint64_t* bar, int row0) {
    mbar_expect_tx(bar, kTileBytes);
    tma_load_3d(dst, tm, bar, 0, row0, bh);
    tma_load_3d(dst + kHalfBytes, tm, bar, 64, row0, bh);
  };
  auto load_kv = [&](const CUtensorMap* tm, int j) {
    int row0, valid; key_tile(u.ks, u.n0, u.j_begin + j, row0, valid);
    const int slot = item % kKvStages;
    mbar_wait(&sm.kv_empty[slot], ((item / kKvStages) & 1) ^ 1);
    load_tile(sm.kv[slot], tm, &sm.kv_full[slot], row0);
    ++item;
  };
  load_tile(sm.q[0], tmq, &sm.q_full[0], u.q_row0[0]);
  load_kv(tmk, 0);
  if (u.n_q == 2) load_tile(sm.q[1], tmq, &sm.q_full[1], u.q_row0[1]);
  load_kv(tmv, 0);
  for (int j = 1; j < u.n_local; ++j) { load_kv(tmk, j); load_kv(tmv, j); }
}

// kLean: the k-loops are not unrolled, so the issuing thread keeps one descriptor pair live
// instead of sixteen (v4 leaves the control warpgroup 32 registers per thread).
template <bool kLean>
__device__ __forceinline__ void v1_mma(SmemV1& sm, uint32_t tmem, const UnitV1& u) {
  constexpr uint32_t idesc_qk = make_idesc_bf16(kBM, kBN, false, false);
  constexpr uint32_t idesc_pv = make_idesc_bf16(kBM, kD, false, true);
  auto kv_wait = [&](int item) {
    const int slot = item % kKvStages;
    mbar_wait(&sm.kv_full[slot], (item / kKvStages) & 1);
    return slot;
  };
  auto qk = [&](int t, int kslot) {                      // S_t = Q_t K^T
    const uint32_t qa = smem_u32(sm.q[t]), ka = smem_u32(sm.kv[kslot]);
    if (kLean) {
#pragma unroll 1
      for (int kk = 0; kk < kD / 16; ++kk) {
        const uint32_t off = (kk >> 2) * kHalfBytes + (kk & 3) * 32;
        mma_ss(tmem + t * 128, make_smem_desc(qa + off, 16, 1024), make_smem_desc(ka + off, 16, 1024),
               idesc_qk, kk > 0);
      }
    } else {
#pragma unroll
      for (int kk = 0; kk < kD / 16; ++kk) {
        const uint32_t off = (kk >> 2) * kHalfBytes + (kk & 3) * 32;
        mma_ss(tmem + t * 128, make_smem_desc(qa + off, 16, 1024), make_smem_desc(ka + off, 16, 1024),
               idesc_qk, kk > 0);
      }
    }
    mma_commit(&sm.s_full[t]);
  };
  auto pv = [&](int t, int vslot, bool acc) {            // O_t (+)= P_t V, P_t from TMEM
    const uint32_t va = smem_u32(sm.kv[vslot]);
    if (kLean) {
#pragma unroll 1
      for (int kk = 0; kk < kBN / 16; ++kk)
        mma_ts(tmem + 256 + t * 128, tmem + t * 128 + kk * 8,
               make_smem_desc(va + kk * 2048, kHalfBytes, 1024), idesc_pv, acc || kk > 0);
    } else {
#pragma unroll
      for (int kk = 0; kk < kBN / 16; ++kk)
        mma_ts(tmem + 256 + t * 128, tmem + t * 128 + kk * 8,
               make_smem_desc(va + kk * 2048, kHalfBytes, 1024), idesc_pv, acc || kk > 0);
    }
  };
  const bool two = u.n_q == 2;
  mbar_wait(&sm.q_full[0], 0);
  int ks = kv_wait(0);
  tc_fence_after();
  qk(0, ks);
  if (two) { mbar_wait(&sm.q_full[1], 0); tc_fence_after(); qk(1, ks); }
  mma_commit(&sm.kv_empty[ks]);
  for (int j = 0; j < u.n_local; ++j) {
    const bool has_next = j + 1 < u.n_local;
    const int vs = kv_wait(2 * j + 1);
    if (has_next) ks = kv_wait(2 * j + 2);
    mbar_wait(&sm.p_full[0], j & 1);
    tc_fence_after();
    pv(0, vs, j > 0);
    if (has_next) qk(0, ks);
    if (two) {
      mbar_wait(&sm.p_full[1], j & 1);
      tc_fence_after();
      pv(1, vs, j > 0);
    }
    mma_commit(&sm.kv_empty[vs]);
    if (has_next) {
      if (two) qk(1, ks);
      mma_commit(&sm.kv_empty[ks]);
    }
  }
  mma_commit(&sm.o_done);
}

// 2^x for x <= 0 on the FMA pipe (Cody-Waite split + degree-3 minimax polynomial, max relative
// error 1.6e-4 << bf16's 3.9e-3): relieves the MUFU.EX2 unit, which needs exactly as many
// cycles per tile as the tensor core.  Two elements per call (packed f32x2 math).
__device__ __forceinline__ void exp2_poly2(uint64_t x2, float& e0, float& e1) {
  float x0, x1;
  upk64(x2, x0, x1);
  x2 = pk64(fmaxf(x0, -126.f), fmaxf(x1, -126.f));
  const uint64_t magic = pk64(12582912.f, 12582912.f), nmagic = pk64(-12582912.f, -12582912.f);
  const uint64_t j2 = add2(x2, magic);                       // round(x) in the low mantissa bits
  const uint64_t f2 = fma2(add2(j2, nmagic), pk64(-1.f, -1.f), x2);     // x - round(x) in [-.5,.5]
  uint64_t p2 = fma2(f2, pk64(0.05360212177f, 0.05360212177f), pk64(0.24237291515f, 0.24237291515f));
  p2 = fma2(p2, f2, pk64(0.69350236654f, 0.69350236654f));
  p2 = fma2(p2, f2, pk64(0.99994814396f, 0.99994814396f));
  float j0, j1, p0, p1;
  upk64(j2, j0, j1); upk64(p2, p0, p1);
  e0 = __int_as_float(__float_as_int(p0) + (__float_as_int(j0) << 23));
  e1 = __int_as_float(__float_as_int(p1) + (__float_as_int(j1) << 23));
}

// Which of the 16 element pairs of a 32-column chunk take the FMA-pipe polynomial instead of
// MUFU.EX2: kPolyEvery = n > 0 -> every n-th pair; -3 -> 3 of every 8 (the two pipes balance at
// ~37 % when the scale / row-sum FMAs are counted); 0 -> none.
template <int kPolyEvery>
__device__ __forceinline__ constexpr bool use_poly(int i) {
  return kPolyEvery > 0 ? (i % (kPolyEvery > 0 ? kPolyEvery : 1)) == kPolyEvery - 1
       : kPolyEvery == -3 ? ((i & 7) == 2 || (i & 7) == 5 || (i & 7) == 7) : false;
}

template <int kPolyEvery>
__device__ __forceinline__ void v1_softmax(SmemV1& sm, uint32_t tmem, const UnitV1& u,
                                           const AttnParams& p, int t, int b, int h) {
  const int warp_q = (threadIdx.x >> 5) & 3, lane = threadIdx.x & 31;
  const int row = warp_q * 32 + lane;
  const uint32_t lane_off = (uint32_t)(warp_q * 32) << 16;
  const uint32_t t_s = tmem + t * 128 + lane_off, t_o = tmem + 256 + t * 128 + lane_off;
  const float scale = p.scale_log2;
  const int q_row0 = t ? u.q_row0[1] : u.q_row0[0], q_rows = t ? u.q_rows[1] : u.q_rows[0];
  float m_used = -INFINITY, l_sum = 0.f;

  // One 32-column chunk of S at a time (two register buffers, the next tcgen05.ld in flight
  // while the current chunk is processed): pass 1 finds the row max, pass 2 re-reads S from
  // TMEM (cheap) and writes P.  Keeps the warpgroup at ~110 live registers instead of holding
  // the whole 128-wide row, which spilled under the 168-register cap of a 384-thread CTA.
  auto mask_chunk = [&](uint32_t* r, int c0, int valid) {
    if (valid < kBN) {                                    // ragged last tile of a key segment
#pragma unroll
      for (int i = 0; i < 32; ++i) if (c0 + i >= valid) r[i] = 0xff800000u;   // -inf
    }
  };
  auto max_chunk = [&](const uint32_t* r, float mx) {
    float m0 = __uint_as_float(r[0]), m1 = __uint_as_float(r[1]);
#pragma unroll
    for (int i = 2; i < 32; i += 2) {
      m0 = fmaxf(m0, __uint_as_float(r[i])); m1 = fmaxf(m1, __uint_as_float(r[i + 1]));
    }
    return fmaxf(mx, fmaxf(m0, m1));
  };

#pragma unroll 1
  for (int j = 0; j < u.n_local; ++j) {
    int row0, valid; key_tile(u.ks, u.n0, u.j_begin + j, row0, valid);
    mbar_wait(&sm.s_full[t], j & 1);
    tc_fence_after();
    uint32_t ra[32], rb[32];
    // ---- pass 1: row max ---------------------------------------------------------------
    float mx = -INFINITY;
    tmem_ld32(t_s, ra);        tmem_wait_ld32(ra);
    tmem_ld32(t_s + 32, rb);   mask_chunk(ra, 0, valid);  mx = max_chunk(ra, mx);  tmem_wait_ld32(rb);
    tmem_ld32(t_s + 64, ra);   mask_chunk(rb, 32, valid); mx = max_chunk(rb, mx);  tmem_wait_ld32(ra);
    tmem_ld32(t_s + 96, rb);   mask_chunk(ra, 64, valid); mx = max_chunk(ra, mx);  tmem_wait_ld32(rb);
    tmem_ld32(t_s, ra);        mask_chunk(rb, 96, valid); mx = max_chunk(rb, mx);  // chunk 0 again, in flight

    const bool grow = (mx - m_used) * scale > kRescaleThreshold;   // true on the first tile
    const float m_new = grow ? mx : m_used;
    if (j > 0 && __any_sync(0xffffffffu, grow)) {
      // PV_t(j-1) has completed: the commit behind s_full[t] covers it.
      const float alpha = grow ? ex2((m_used - m_new) * scale) : 1.f;
      l_sum *= alpha;
      tmem_wait_ld32(ra);
#pragma unroll
      for (int c = 0; c < kD; c += 32) {
        tmem_ld32(t_o + c, rb);
        tmem_wait_ld32(rb);
#pragma unroll
        for (int i = 0; i < 32; ++i) rb[i] = __float_as_uint(__uint_as_float(rb[i]) * alpha);
        tmem_st32(t_o + c, rb);
      }
    }
    m_used = m_new;
    const float mb = m_used * scale;
    const uint64_t scale2 = pk64(scale, scale), nmb2 = pk64(-mb, -mb);
    uint64_t rs2 = pk64(0.f, 0.f);
    // ---- pass 2: P = exp2(S*scale - m), bf16, written over S columns [0,64) ----------------
    auto exp_chunk = [&](uint32_t* r, int c0) {
      mask_chunk(r, c0, valid);
      uint32_t pk[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const uint64_t x2 = fma2(pk64(__uint_as_float(r[2 * i]), __uint_as_float(r[2 * i + 1])), scale2, nmb2);
        float e0, e1;
        if (kPolyEvery > 0 && (i % (kPolyEvery > 0 ? kPolyEvery : 1)) == kPolyEvery - 1) {
          exp2_poly2(x2, e0, e1);
        } else {
          float x0, x1;
          upk64(x2, x0, x1);
          e0 = ex2(x0); e1 = ex2(x1);
        }
        rs2 = add2(rs2, pk64(e0, e1));
        pk[i] = pack2(e0, e1);
      }
      tmem_st16(t_s + (c0 >> 1), pk);
    };
    tmem_wait_ld32(ra);
    tmem_ld32(t_s + 32, rb);   exp_chunk(ra, 0);    tmem_wait_ld32(rb);
    tmem_ld32(t_s + 64, ra);   exp_chunk(rb, 32);   tmem_wait_ld32(ra);
    tmem_ld32(t_s + 96, rb);   exp_chunk(ra, 64);   tmem_wait_ld32(rb);
    exp_chunk(rb, 96);
    {
      float r0, r1;
      upk64(rs2, r0, r1);
      l_sum += r0 + r1;
    }
    tmem_wait_st();
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(&sm.p_full[t]);
  }

  // ---- epilogue: O_t / l -> bf16 "B L (H D)" ---------------------------------------
  mbar_wait(&sm.o_done, 0);
  tc_fence_after();
  const float inv_l = 1.f / l_sum;
#pragma unroll
  for (int c = 0; c < kD; c += 32) {
    uint32_t r[32];
    tmem_ld32(t_o + c, r);
    tmem_wait_ld();
    if (row < q_rows) {
      float o[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) o[i] = __uint_as_float(r[i]);
      store_out_row(p, b, h, q_row0 + row, o, inv_l, c, 32);
    }
  }
}

// ---------------------------------------------------------------------------------------
// v3 softmax: BOTH softmax warpgroups work on the SAME tile, split by key columns (warpgroup
// `half` owns S columns [64*half, 64*half+64)), then move to the other tile together.
// Why: MUFU.EX2 is quarter rate — a warp that has just issued one cannot issue its next MUFU
// for 8 cycles, and ptxas clusters the MUFUs of a chunk, so a lone warp per SM sub-partition
// (v1/v2: one warpgroup per tile, the two alternating) idles through those slots (ncu: 592
// instructions took ~2000 cycles).  With two warps of the same tile co-resident on each
// sub-partition one warp's FMA/pack work fills the other's MUFU shadow.
// The two threads that share a row exchange their partial row max through shared memory
// (named barrier, 256 threads); that barrier also orders every S read of the tile before
// any P write, which matters because P (columns [0,64)) overlays the columns warpgroup 0 reads.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ void softmax_bar_sync(int id) {
  asm volatile("bar.sync %0, 256;" ::"r"(id) : "memory");
}

// kTile = -1: v3, the 256 softmax threads walk tile 0 then tile 1 of every key step.
// kTile = 0/1: v4, this 256-thread pair owns tile kTile only; the other pair (a second set of
// two warpgroups) owns the other tile and runs concurrently, staggered by the MMA pipeline, so
// each SM sub-partition holds four softmax warps in two different phases: one pair's
// TMEM-load / max / exchange / pack / TMEM-store latency hides under the other's MUFU stream.
// kOpt (optimistic max): after the first key tile the exp2 stream starts at once with the row
// max already in use and the exchange of this tile's max moves behind it (same single barrier
// per tile, which still orders every S read before any P write).  Only when the exchanged max
// turns out to exceed the one in use by more than the lazy-rescale threshold — rare after the
// first tiles — is the tile redone from the S values still in registers, exactly as the plain
// path would have done it: results are bit-identical.
template <int kPolyEvery, int kTile, bool kOpt = false>
__device__ __forceinline__ void v3_softmax(SmemV1& sm, uint32_t tmem, const UnitV1& u,
                                           const AttnParams& p, int half, int b, int h) {
  constexpr int kPair = kTile < 0 ? 0 : kTile;
  const int warp_q = (threadIdx.x >> 5) & 3, lane = threadIdx.x & 31;
  const int row = warp_q * 32 + lane;
  const uint32_t lane_off = (uint32_t)(warp_q * 32) << 16;
  const float scale = p.scale_log2;
  const int c0 = half * 64;
  float m_used[2] = {-INFINITY, -INFINITY}, l_sum[2] = {0.f, 0.f};
  uint32_t use = 0;                                       // exchange-buffer parity

  auto mask_chunk = [&](uint32_t* r, int cbase, int valid) {
    if (valid < kBN) {                                    // ragged last tile of a key segment
#pragma unroll
      for (int i = 0; i < 32; ++i) if (cbase + i >= valid) r[i] = 0xff800000u;   // -inf
    }
  };
  auto max_chunk = [&](const uint32_t* r, float mx) {
    float m0 = __uint_as_float(r[0]), m1 = __uint_as_float(r[1]);
#pragma unroll
    for (int i = 2; i < 32; i += 2) {
      m0 = fmaxf(m0, __uint_as_float(r[i])); m1 = fmaxf(m1, __uint_as_float(r[i + 1]));
    }
    return fmaxf(mx, fmaxf(m0, m1));
  };
  auto exchange = [&](float mine, bool take_max) {        // combine with the row's other thread
    float* xb = &sm.xchg[kPair][use & 1][0][0];
    xb[half * kBM + row] = mine;
    softmax_bar_sync(1 + kPair);
    const float other = xb[(half ^ 1) * kBM + row];
    ++use;
    return take_max ? fmaxf(mine, other) : mine + other;
  };

#pragma unroll 1
  for (int j = 0; j < u.n_local; ++j) {
    int row0, valid; key_tile(u.ks, u.n0, u.j_begin + j, row0, valid);
#pragma unroll
    for (int t = 0; t < 2; ++t) {
      if ((kTile < 0 || t == kTile) && t < u.n_q) {
        const uint32_t t_s = tmem + t * 128 + lane_off, t_o = tmem + 256 + t * 128 + lane_off;
        mbar_wait(&sm.s_full[t], j & 1);
        tc_fence_after();
        uint32_t ra[32], rb[32];
        tmem_ld32(t_s + c0, ra);
        tmem_ld32(t_s + c0 + 32, rb);
        tmem_wait_ld32(ra);
        tmem_wait_ld32(rb);
        mask_chunk(ra, c0, valid);
        mask_chunk(rb, c0 + 32, valid);
        const float m_loc = max_chunk(rb, max_chunk(ra, -INFINITY));
        // P = exp2(S * scale - m * scale) for one 32-column chunk -> 16 packed bf16 pairs
        auto exp_regs = [&](const uint32_t* r, float m_ref, uint32_t* pk, uint64_t& rs2) {
          const float mb = m_ref * scale;
          const uint64_t scale2 = pk64(scale, scale), nmb2 = pk64(-mb, -mb);
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const uint64_t x2 = fma2(pk64(__uint_as_float(r[2 * i]), __uint_as_float(r[2 * i + 1])), scale2, nmb2);
            float e0, e1;
            if (use_poly<kPolyEvery>(i)) {
              exp2_poly2(x2, e0, e1);
            } else {
              float x0, x1;
              upk64(x2, x0, x1);
              e0 = ex2(x0); e1 = ex2(x1);
            }
            rs2 = add2(rs2, pk64(e0, e1));
            pk[i] = pack2(e0, e1);
          }
        };
        auto rescale_o = [&](float alpha) {
          // rare path: 8 columns at a time so it adds nothing to the loop's register budget
#pragma unroll 1
          for (int c = 0; c < 64; c += 8) {
            uint32_t rc[8];
            tmem_ld8(t_o + c0 + c, rc);
            tmem_wait_ld8(rc);
#pragma unroll
            for (int i = 0; i < 8; ++i) rc[i] = __float_as_uint(__uint_as_float(rc[i]) * alpha);
            tmem_st8(t_o + c0 + c, rc);
          }
        };
        uint32_t pka[16], pkb[16];
        uint64_t rs2 = pk64(0.f, 0.f);
        if (kOpt && j > 0) {
          float* xb = &sm.xchg[kPair][use & 1][0][0];
          xb[half * kBM + row] = m_loc;
          exp_regs(ra, m_used[t], pka, rs2);              // optimistic: the max in use still holds
          exp_regs(rb, m_used[t], pkb, rs2);
          softmax_bar_sync(1 + kPair);
          const float mx = fmaxf(m_loc, xb[(half ^ 1) * kBM + row]);
          ++use;
          const bool grow = (mx - m_used[t]) * scale > kRescaleThreshold;
          if (__any_sync(0xffffffffu, grow)) {            // rare: redo the tile with the new max
            const float m_new = grow ? mx : m_used[t];
            const float alpha = grow ? ex2((m_used[t] - m_new) * scale) : 1.f;
            l_sum[t] *= alpha;
            rescale_o(alpha);                             // PV_t(j-1) is complete (see s_full)
            m_used[t] = m_new;
            rs2 = pk64(0.f, 0.f);
            exp_regs(ra, m_new, pka, rs2);
            exp_regs(rb, m_new, pkb, rs2);
          }
        } else {
          const float mx = exchange(m_loc, true);
          const bool grow = (mx - m_used[t]) * scale > kRescaleThreshold;   // true on the first tile
          const float m_new = grow ? mx : m_used[t];
          if (j > 0 && __any_sync(0xffffffffu, grow)) {
            // PV_t(j-1) has completed: the commit behind s_full[t] covers it.
            const float alpha = grow ? ex2((m_used[t] - m_new) * scale) : 1.f;
            l_sum[t] *= alpha;
            rescale_o(alpha);
          }
          m_used[t] = m_new;
          exp_regs(ra, m_new, pka, rs2);
          exp_regs(rb, m_new, pkb, rs2);
        }
        tmem_st16(t_s + half * 32, pka);                  // P_t over S_t columns [0,64)
        tmem_st16(t_s + half * 32 + 16, pkb);
        float r0, r1;
        upk64(rs2, r0, r1);
        l_sum[t] += r0 + r1;
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.p_full[t]);
      }
    }
  }

  // ---- epilogue: O_t / l -> bf16 "B L (H D)"; this thread owns columns [c0, c0+64) -------
  mbar_wait(&sm.o_done, 0);
  tc_fence_after();
  if (u.part_slot < 0) {
#pragma unroll
    for (int t = 0; t < 2; ++t) {
      if ((kTile < 0 || t == kTile) && t < u.n_q) {
        const uint32_t t_o = tmem + 256 + t * 128 + lane_off;
        const float inv_l = 1.f / exchange(l_sum[t], false);
        const int q_row0 = t ? u.q_row0[1] : u.q_row0[0], q_rows = t ? u.q_rows[1] : u.q_rows[0];
#pragma unroll
        for (int c = 0; c < 64; c += 32) {
          uint32_t r[32];
          tmem_ld32(t_o + c0 + c, r);
          tmem_wait_ld32(r);
          if (row < q_rows) {
            float o[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __uint_as_float(r[i]);
            store_out_row(p, b, h, q_row0 + row, o, inv_l, c0 + c, 32);
          }
        }
      }
    }
    return;
  }

  // ---- stream-K piece: this CTA saw only key tiles [j_begin, j_begin + n_local) of its unit.
  // Write (O unnormalised, row max, row sum) to the piece's slot; the piece that arrives last
  // (atomic ticket, no waiting) merges all pieces of the unit in a fixed order and stores.
  if (kTile >= 0) return;                                 // only the v3 layout takes part
  const int n_main = p.n_main_tiles;
  {
    float* slot = p.ws_part + (size_t)u.part_slot * kPartFloats;
#pragma unroll
    for (int t = 0; t < 2; ++t) {
      if (t < u.n_q) {
        const uint32_t t_o = tmem + 256 + t * 128 + lane_off;
        float* pt = slot + t * kPartTile;
        const float l_tot = exchange(l_sum[t], false);
        if (half == 0) { pt[kBM * kD + row] = m_used[t]; pt[kBM * kD + kBM + row] = l_tot; }
#pragma unroll
        for (int c = 0; c < 64; c += 32) {
          uint32_t r[32];
          tmem_ld32(t_o + c0 + c, r);
          tmem_wait_ld32(r);
#pragma unroll
          for (int i = 0; i < 32; i += 4)
            *reinterpret_cast<uint4*>(pt + (((c0 + c + i) >> 2) * kBM + row) * 4) =
                make_uint4(r[i], r[i + 1], r[i + 2], r[i + 3]);   // [col/4][row][4]: a warp stores 512 B runs
        }
      }
    }
  }
  // pieces of this unit: ranges i_first.. that overlap key steps [a*n, (a+1)*n)
  const int64_t u_lo = (int64_t)u.sk_unit * n_main, u_hi = u_lo + n_main;
  const int64_t total = (int64_t)p.sk_units * n_main;
  const int i_first = (int)((u_lo * p.sk_sms) / total);
  int n_pieces = 0;
  for (int i = max(0, i_first - 1); i < p.sk_sms; ++i) {
    int64_t lo, hi;
    sk_range(p, n_main, i, lo, hi);
    if (lo >= u_hi) break;
    if (hi > u_lo && hi > lo) ++n_pieces;
  }
  __threadfence();
  softmax_bar_sync(1);
  float* xflag = &sm.xchg[0][0][0][0];
  if (threadIdx.x == 128) {                               // first softmax thread
    const int ticket = atomicAdd(p.ws_count + u.sk_unit, 1);
    xflag[0] = ticket == n_pieces - 1 ? 1.f : 0.f;
  }
  softmax_bar_sync(1);
  if (xflag[0] == 0.f) return;
  __threadfence();
#pragma unroll 1
  for (int t = 0; t < 2; ++t) {
    if (t >= u.n_q) break;
    const int q_row0 = t ? u.q_row0[1] : u.q_row0[0], q_rows = t ? u.q_rows[1] : u.q_rows[0];
    // pass 1: the row max over the pieces
    float m_all = -INFINITY;
    for (int i = max(0, i_first - 1); i < p.sk_sms; ++i) {
      int64_t lo, hi;
      sk_range(p, n_main, i, lo, hi);
      if (lo >= u_hi) break;
      if (!(hi > u_lo && hi > lo)) continue;
      const int k = (lo / n_main == u.sk_unit) ? 0 : 1;
      const float* pt = p.ws_part + (size_t)(2 * i + k) * kPartFloats + t * kPartTile;
      m_all = fmaxf(m_all, __ldcg(pt + kBM * kD + row));
    }
    // pass 2: O = sum_s O_s * 2^((m_s - m_all) * scale), l likewise; 32 columns at a time
#pragma unroll 1
    for (int c = 0; c < 64; c += 32) {
      float acc[32], l_all = 0.f;
#pragma unroll
      for (int e = 0; e < 32; ++e) acc[e] = 0.f;
      for (int i = max(0, i_first - 1); i < p.sk_sms; ++i) {
        int64_t lo, hi;
        sk_range(p, n_main, i, lo, hi);
        if (lo >= u_hi) break;
        if (!(hi > u_lo && hi > lo)) continue;
        const int k = (lo / n_main == u.sk_unit) ? 0 : 1;
        const float* pt = p.ws_part + (size_t)(2 * i + k) * kPartFloats + t * kPartTile;
        const float f = ex2((__ldcg(pt + kBM * kD + row) - m_all) * scale);
        l_all += f * __ldcg(pt + kBM * kD + kBM + row);
#pragma unroll
        for (int e = 0; e < 32; e += 4) {
          const float4 o4 = __ldcg(reinterpret_cast<const float4*>(pt + (((c0 + c + e) >> 2) * kBM + row) * 4));
          acc[e] += f * o4.x; acc[e + 1] += f * o4.y; acc[e + 2] += f * o4.z; acc[e + 3] += f * o4.w;
        }
      }
      if (row < q_rows) store_out_row(p, b, h, q_row0 + row, acc, 1.f / l_all, c0 + c, 32);
    }
  }
  if (threadIdx.x == 128) p.ws_count[u.sk_unit] = 0;      // ready for the next launch
}

// kMode 0: v1/v2 (one softmax warpgroup per tile), 1: v3 (two warpgroups, one tile at a time),
// 2: v4 (four warpgroups = two column-half pairs, one pair per tile, 640 threads; MMA issue loops
// rolled — unrolled they spill at 32 registers and measured 5-20 % slower).
// 3: v3 with the optimistic row max (see v3_softmax).
template <int kPolyEvery, int kMode>
__global__ void __launch_bounds__(kMode == 2 ? 640 : 384, 1)
attn_v1_kernel(const __grid_constant__ CUtensorMap tmq, const __grid_constant__ CUtensorMap tmk,
               const __grid_constant__ CUtensorMap tmv, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  SmemV1& sm = *reinterpret_cast<SmemV1*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const UnitV1 u = decode_unit(p, blockIdx.x);
  if (u.n_local == 0) return;                       // empty stream-K piece (whole CTA, before any barrier)
  const int bh = u.bh, b = bh / p.H, h = bh - b * p.H;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmq); prefetch_tmap(&tmk); prefetch_tmap(&tmv);
    mbar_init(&sm.q_full[0], 1); mbar_init(&sm.q_full[1], 1);
    for (int i = 0; i < kKvStages; ++i) { mbar_init(&sm.kv_full[i], 1); mbar_init(&sm.kv_empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&sm.s_full[i], 1); mbar_init(&sm.p_full[i], kMode ? 8 : 4); }
    mbar_init(&sm.o_done, 1);
    fence_barrier_init();
  }
  if (warp == 2) { tmem_alloc(&sm.tmem_base, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  if (warp < 4) {
    if (kMode == 2) setmaxnreg_dec<32>(); else setmaxnreg_dec<96>();
    if (warp == 0) {
      if (elect_one()) v1_producer(sm, &tmq, &tmk, &tmv, u, bh);
    } else if (warp == 1) {
      if (elect_one()) v1_mma<kMode == 2>(sm, tmem, u);
    }
  } else if (kMode == 2) {
    // setmaxnreg pool = what the CTA was launched with, 640 x 96 = 61440 = 4 x 32 x 32 + 16 x 32 x 112
    // (asking for more blocks forever)
    setmaxnreg_inc<112>();
    const int pair = (warp - 4) >> 3, half = ((warp - 4) >> 2) & 1;
    if (pair == 0) v3_softmax<kPolyEvery, 0>(sm, tmem, u, p, half, b, h);
    else v3_softmax<kPolyEvery, 1>(sm, tmem, u, p, half, b, h);
  } else {
    setmaxnreg_inc<200>();
    const int t = (warp >> 2) - 1;
    if (kMode == 1) v3_softmax<kPolyEvery, -1>(sm, tmem, u, p, t, b, h);
    else if (kMode == 3) v3_softmax<kPolyEvery, -1, true>(sm, tmem, u, p, t, b, h);
    else if (t < u.n_q) v1_softmax<kPolyEvery>(sm, tmem, u, p, t, b, h);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem, 512);
}

// =============================================================================
// Persistent form of v3 (impl 14, EXPERIMENTAL: opt-in through ca_debug_set_attention_impl).
// One CTA per SM walks the flat unit list [concept | whole | stream-K pieces] with a static
// stride; TMEM is allocated and the barriers are initialised once.  All mbarrier parities come
// from running counters (K/V ring items, per-tile S/P completions, units) instead of `j & 1`.
// What this buys: the next unit's Q/K/V loads and its first QK^T run under the
// current unit's epilogue, and the per-CTA start-up (TMEM alloc, barrier init, tensor-map
// prefetch, first TMA round trip) is paid once per SM instead of once per unit.
// Ordering that makes it safe without an "O free" barrier: the first PV of unit n+1 needs
// p_full, which all eight softmax warps arrive on only after their epilogue of unit n has read O
// out of TMEM; the next unit's QK^T overwrites S/P only behind the last PV in the tensor pipe's
// issue order.  One new barrier, q_empty (MMA warp -> producer): the unit's last QK^T, the last
// reader of sm.q, has been committed.
// =============================================================================
struct SmemP : SmemV1 {
  uint64_t q_empty;
  float sk_flag[4];
};

__device__ __forceinline__ void pz_producer(SmemP& sm, const CUtensorMap* tmq, const CUtensorMap* tmk,
                                            const CUtensorMap* tmv, const AttnParams& p, int total) {
  int item = 0, n_unit = 0;
  for (int cta = blockIdx.x; cta < total; cta += gridDim.x) {
    const UnitV1 u = decode_unit(p, cta);
    if (u.n_local == 0) continue;
    const int bh = u.bh;
    auto load_tile = [&](uint8_t* dst, const CUtensorMap* tm, uint64_t* bar, int row0) {
      mbar_expect_tx(bar, kTileBytes);
      tma_load_3d(dst, tm, bar, 0, row0, bh);
      tma_load_3d(dst + kHalfBytes, tm, bar, 64, row0, bh);
    };
    auto load_kv = [&](const CUtensorMap* tm, int j) {
      int row0, valid; key_tile(u.ks, u.n0, u.j_begin + j, row0, valid);
      const int slot = item % kKvStages;
      mbar_wait(&sm.kv_empty[slot], ((item / kKvStages) & 1) ^ 1);
      load_tile(sm.kv[slot], tm, &sm.kv_full[slot], row0);
      ++item;
    };
    if (n_unit > 0) mbar_wait(&sm.q_empty, (n_unit - 1) & 1);   // previous unit's QK^Ts are done with sm.q
    load_tile(sm.q[0], tmq, &sm.q_full[0], u.q_row0[0]);
    load_kv(tmk, 0);
    if (u.n_q == 2) load_tile(sm.q[1], tmq, &sm.q_full[1], u.q_row0[1]);
    load_kv(tmv, 0);
    for (int j = 1; j < u.n_local; ++j) { load_kv(tmk, j); load_kv(tmv, j); }
    ++n_unit;
  }
}

__device__ __forceinline__ void pz_mma(SmemP& sm, uint32_t tmem, const AttnParams& p, int total) {
  constexpr uint32_t idesc_qk = make_idesc_bf16(kBM, kBN, false, false);
  constexpr uint32_t idesc_pv = make_idesc_bf16(kBM, kD, false, true);
  int item = 0;                       // K/V ring items consumed before this unit
  int n_qf[2] = {0, 0};               // q_full[t] completions consumed
  int n_pf[2] = {0, 0};               // p_full[t] completions consumed
  auto kv_wait = [&](int it) {
    const int slot = it % kKvStages;
    mbar_wait(&sm.kv_full[slot], (it / kKvStages) & 1);
    return slot;
  };
  auto qk = [&](int t, int kslot) {                      // S_t = Q_t K^T
    const uint32_t qa = smem_u32(sm.q[t]), ka = smem_u32(sm.kv[kslot]);
#pragma unroll
    for (int kk = 0; kk < kD / 16; ++kk) {
      const uint32_t off = (kk >> 2) * kHalfBytes + (kk & 3) * 32;
      mma_ss(tmem + t * 128, make_smem_desc(qa + off, 16, 1024), make_smem_desc(ka + off, 16, 1024),
             idesc_qk, kk > 0);
    }
    mma_commit(&sm.s_full[t]);
  };
  auto pv = [&](int t, int vslot, bool acc) {            // O_t (+)= P_t V, P_t from TMEM
    const uint32_t va = smem_u32(sm.kv[vslot]);
#pragma unroll
    for (int kk = 0; kk < kBN / 16; ++kk)
      mma_ts(tmem + 256 + t * 128, tmem + t * 128 + kk * 8,
             make_smem_desc(va + kk * 2048, kHalfBytes, 1024), idesc_pv, acc || kk > 0);
  };
  for (int cta = blockIdx.x; cta < total; cta += gridDim.x) {
    const UnitV1 u = decode_unit(p, cta);
    if (u.n_local == 0) continue;
    const bool two = u.n_q == 2;
    mbar_wait(&sm.q_full[0], n_qf[0]++ & 1);
    int ks = kv_wait(item);
    tc_fence_after();
    qk(0, ks);
    if (two) { mbar_wait(&sm.q_full[1], n_qf[1]++ & 1); tc_fence_after(); qk(1, ks); }
    mma_commit(&sm.kv_empty[ks]);
    if (u.n_local == 1) mma_commit(&sm.q_empty);
    for (int j = 0; j < u.n_local; ++j) {
      const bool has_next = j + 1 < u.n_local;
      const int vs = kv_wait(item + 2 * j + 1);
      if (has_next) ks = kv_wait(item + 2 * j + 2);
      mbar_wait(&sm.p_full[0], n_pf[0]++ & 1);
      tc_fence_after();
      pv(0, vs, j > 0);
      if (has_next) qk(0, ks);
      if (two) {
        mbar_wait(&sm.p_full[1], n_pf[1]++ & 1);
        tc_fence_after();
        pv(1, vs, j > 0);
      }
      mma_commit(&sm.kv_empty[vs]);
      if (has_next) {
        if (two) qk(1, ks);
        mma_commit(&sm.kv_empty[ks]);
        if (j + 2 == u.n_local) mma_commit(&sm.q_empty);   // that was the unit's last QK^T
      }
    }
    mma_commit(&sm.o_done);
    item += 2 * u.n_local;
  }
}

template <int kPolyEvery>
__device__ __forceinline__ void pz_softmax(SmemP& sm, uint32_t tmem, const AttnParams& p, int half,
                                           int total) {
  const int warp_q = (threadIdx.x >> 5) & 3, lane = threadIdx.x & 31;
  const int row = warp_q * 32 + lane;
  const uint32_t lane_off = (uint32_t)(warp_q * 32) << 16;
  const float scale = p.scale_log2;
  const int c0 = half * 64;
  uint32_t use = 0;                                       // exchange-buffer parity (runs across units)
  int n_sf[2] = {0, 0};                                   // s_full[t] completions consumed
  int n_unit = 0;                                         // o_done completions consumed

  auto mask_chunk = [&](uint32_t* r, int cbase, int valid) {
    if (valid < kBN) {
#pragma unroll
      for (int i = 0; i < 32; ++i) if (cbase + i >= valid) r[i] = 0xff800000u;   // -inf
    }
  };
  auto max_chunk = [&](const uint32_t* r, float mx) {
    float m0 = __uint_as_float(r[0]), m1 = __uint_as_float(r[1]);
#pragma unroll
    for (int i = 2; i < 32; i += 2) {
      m0 = fmaxf(m0, __uint_as_float(r[i])); m1 = fmaxf(m1, __uint_as_float(r[i + 1]));
    }
    return fmaxf(mx, fmaxf(m0, m1));
  };
  auto exchange = [&](float mine, bool take_max) {        // combine with the row's other thread
    float* xb = &sm.xchg[0][use & 1][0][0];
    xb[half * kBM + row] = mine;
    softmax_bar_sync(1);
    const float other = xb[(half ^ 1) * kBM + row];
    ++use;
    return take_max ? fmaxf(mine, other) : mine + other;
  };

#pragma unroll 1
  for (int cta = blockIdx.x; cta < total; cta += gridDim.x) {
    const UnitV1 u = decode_unit(p, cta);
    if (u.n_local == 0) continue;
    const int b = u.bh / p.H, h = u.bh - b * p.H;
    float m_used[2] = {-INFINITY, -INFINITY}, l_sum[2] = {0.f, 0.f};

#pragma unroll 1
    for (int j = 0; j < u.n_local; ++j) {
      int row0, valid; key_tile(u.ks, u.n0, u.j_begin + j, row0, valid);
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        if (t < u.n_q) {
          const uint32_t t_s = tmem + t * 128 + lane_off, t_o = tmem + 256 + t * 128 + lane_off;
          mbar_wait(&sm.s_full[t], n_sf[t]++ & 1);
          tc_fence_after();
          uint32_t ra[32], rb[32];
          tmem_ld32(t_s + c0, ra);
          tmem_ld32(t_s + c0 + 32, rb);
          tmem_wait_ld32(ra);
          tmem_wait_ld32(rb);
          mask_chunk(ra, c0, valid);
          mask_chunk(rb, c0 + 32, valid);
          const float m_loc = max_chunk(rb, max_chunk(ra, -INFINITY));
          auto exp_regs = [&](const uint32_t* r, float m_ref, uint32_t* pk, uint64_t& rs2) {
            const float mb = m_ref * scale;
            const uint64_t scale2 = pk64(scale, scale), nmb2 = pk64(-mb, -mb);
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const uint64_t x2 = fma2(pk64(__uint_as_float(r[2 * i]), __uint_as_float(r[2 * i + 1])), scale2, nmb2);
              float e0, e1;
              if (use_poly<kPolyEvery>(i)) {
                exp2_poly2(x2, e0, e1);
              } else {
                float x0, x1;
                upk64(x2, x0, x1);
                e0 = ex2(x0); e1 = ex2(x1);
              }
              rs2 = add2(rs2, pk64(e0, e1));
              pk[i] = pack2(e0, e1);
            }
          };
          auto rescale_o = [&](float alpha) {
#pragma unroll 1
            for (int c = 0; c < 64; c += 8) {
              uint32_t rc[8];
              tmem_ld8(t_o + c0 + c, rc);
              tmem_wait_ld8(rc);
#pragma unroll
              for (int i = 0; i < 8; ++i) rc[i] = __float_as_uint(__uint_as_float(rc[i]) * alpha);
              tmem_st8(t_o + c0 + c, rc);
            }
          };
          uint32_t pka[16], pkb[16];
          uint64_t rs2 = pk64(0.f, 0.f);
          const float mx = exchange(m_loc, true);
          const bool grow = (mx - m_used[t]) * scale > kRescaleThreshold;   // true on the first tile
          const float m_new = grow ? mx : m_used[t];
          if (j > 0 && __any_sync(0xffffffffu, grow)) {
            const float alpha = grow ? ex2((m_used[t] - m_new) * scale) : 1.f;
            l_sum[t] *= alpha;
            rescale_o(alpha);
          }
          m_used[t] = m_new;
          exp_regs(ra, m_new, pka, rs2);
          exp_regs(rb, m_new, pkb, rs2);
          tmem_st16(t_s + half * 32, pka);
          tmem_st16(t_s + half * 32 + 16, pkb);
          float r0, r1;
          upk64(rs2, r0, r1);
          l_sum[t] += r0 + r1;
          tmem_wait_st();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&sm.p_full[t]);
        }
      }
    }

    // ---- epilogue of the unit -----------------------------------------------------------
    mbar_wait(&sm.o_done, n_unit++ & 1);
    tc_fence_after();
    if (u.part_slot < 0) {
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        if (t < u.n_q) {
          const uint32_t t_o = tmem + 256 + t * 128 + lane_off;
          const float inv_l = 1.f / exchange(l_sum[t], false);
          const int q_row0 = t ? u.q_row0[1] : u.q_row0[0], q_rows = t ? u.q_rows[1] : u.q_rows[0];
#pragma unroll
          for (int c = 0; c < 64; c += 32) {
            uint32_t r[32];
            tmem_ld32(t_o + c0 + c, r);
            tmem_wait_ld32(r);
            if (row < q_rows) {
              float o[32];
#pragma unroll
              for (int i = 0; i < 32; ++i) o[i] = __uint_as_float(r[i]);
              store_out_row(p, b, h, q_row0 + row, o, inv_l, c0 + c, 32);
            }
          }
        }
      }
      continue;
    }

    // ---- stream-K piece: partial to its slot, last arriver merges (as in v3_softmax) -------
    const int n_main = p.n_main_tiles;
    {
      float* slot = p.ws_part + (size_t)u.part_slot * kPartFloats;
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        if (t < u.n_q) {
          const uint32_t t_o = tmem + 256 + t * 128 + lane_off;
          float* pt = slot + t * kPartTile;
          const float l_tot = exchange(l_sum[t], false);
          if (half == 0) { pt[kBM * kD + row] = m_used[t]; pt[kBM * kD + kBM + row] = l_tot; }
#pragma unroll
          for (int c = 0; c < 64; c += 32) {
            uint32_t r[32];
            tmem_ld32(t_o + c0 + c, r);
            tmem_wait_ld32(r);
#pragma unroll
            for (int i = 0; i < 32; i += 4)
              *reinterpret_cast<uint4*>(pt + (((c0 + c + i) >> 2) * kBM + row) * 4) =
                  make_uint4(r[i], r[i + 1], r[i + 2], r[i + 3]);
          }
        }
      }
    }
    const int64_t u_lo = (int64_t)u.sk_unit * n_main, u_hi = u_lo + n_main;
    const int64_t total_steps = (int64_t)p.sk_units * n_main;
    const int i_first = (int)((u_lo * p.sk_sms) / total_steps);
    int n_pieces = 0;
    for (int i = max(0, i_first - 1); i < p.sk_sms; ++i) {
      int64_t lo, hi;
      sk_range(p, n_main, i, lo, hi);
      if (lo >= u_hi) break;
      if (hi > u_lo && hi > lo) ++n_pieces;
    }
    __threadfence();
    softmax_bar_sync(1);
    if (threadIdx.x == 128) {                               // first softmax thread
      const int ticket = atomicAdd(p.ws_count + u.sk_unit, 1);
      sm.sk_flag[0] = ticket == n_pieces - 1 ? 1.f : 0.f;
    }
    softmax_bar_sync(1);
    const bool merge = sm.sk_flag[0] != 0.f;
    softmax_bar_sync(1);                                    // flag read by all before the next piece rewrites it
    if (!merge) continue;
    __threadfence();
#pragma unroll 1
    for (int t = 0; t < 2; ++t) {
      if (t >= u.n_q) break;
      const int q_row0 = t ? u.q_row0[1] : u.q_row0[0], q_rows = t ? u.q_rows[1] : u.q_rows[0];
      float m_all = -INFINITY;
      for (int i = max(0, i_first - 1); i < p.sk_sms; ++i) {
        int64_t lo, hi;
        sk_range(p, n_main, i, lo, hi);
        if (lo >= u_hi) break;
        if (!(hi > u_lo && hi > lo)) continue;
        const int k = (lo / n_main == u.sk_unit) ? 0 : 1;
        const float* pt = p.ws_part + (size_t)(2 * i + k) * kPartFloats + t * kPartTile;
        m_all = fmaxf(m_all, __ldcg(pt + kBM * kD + row));
      }
#pragma unroll 1
      for (int c = 0; c < 64; c += 32) {
        float acc[32], l_all = 0.f;
#pragma unroll
        for (int e = 0; e < 32; ++e) acc[e] = 0.f;
        for (int i = max(0, i_first - 1); i < p.sk_sms; ++i) {
          int64_t lo, hi;
          sk_range(p, n_main, i, lo, hi);
          if (lo >= u_hi) break;
          if (!(hi > u_lo && hi > lo)) continue;
          const int k = (lo / n_main == u.sk_unit) ? 0 : 1;
          const float* pt = p.ws_part + (size_t)(2 * i + k) * kPartFloats + t * kPartTile;
          const float f = ex2((__ldcg(pt + kBM * kD + row) - m_all) * scale);
          l_all += f * __ldcg(pt + kBM * kD + kBM + row);
#pragma unroll
          for (int e = 0; e < 32; e += 4) {
            const float4 o4 = __ldcg(reinterpret_cast<const float4*>(pt + (((c0 + c + e) >> 2) * kBM + row) * 4));
            acc[e] += f * o4.x; acc[e + 1] += f * o4.y; acc[e + 2] += f * o4.z; acc[e + 3] += f * o4.w;
          }
        }
        if (row < q_rows) store_out_row(p, b, h, q_row0 + row, acc, 1.f / l_all, c0 + c, 32);
      }
    }
    if (threadIdx.x == 128) p.ws_count[u.sk_unit] = 0;      // ready for the next launch
  }
}

template <int kPolyEvery>
__global__ void __launch_bounds__(384, 1)
attn_persist_kernel(const __grid_constant__ CUtensorMap tmq, const __grid_constant__ CUtensorMap tmk,
                    const __grid_constant__ CUtensorMap tmv, const AttnParams p, const int total) {
  extern __shared__ uint8_t smem_raw[];
  SmemP& sm = *reinterpret_cast<SmemP*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  if (threadIdx.x == 0) {
    prefetch_tmap(&tmq); prefetch_tmap(&tmk); prefetch_tmap(&tmv);
    mbar_init(&sm.q_full[0], 1); mbar_init(&sm.q_full[1], 1);
    for (int i = 0; i < kKvStages; ++i) { mbar_init(&sm.kv_full[i], 1); mbar_init(&sm.kv_empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&sm.s_full[i], 1); mbar_init(&sm.p_full[i], 8); }
    mbar_init(&sm.o_done, 1);
    mbar_init(&sm.q_empty, 1);
    fence_barrier_init();
  }
  if (warp == 2) { tmem_alloc(&sm.tmem_base, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;
  if (warp < 4) {
    setmaxnreg_dec<96>();
    if (warp == 0) {
      if (elect_one()) pz_producer(sm, &tmq, &tmk, &tmv, p, total);
    } else if (warp == 1) {
      if (elect_one()) pz_mma(sm, tmem, p, total);
    }
  } else {
    setmaxnreg_inc<200>();
    pz_softmax<kPolyEvery>(sm, tmem, p, (warp >> 2) - 1, total);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem, 512);
}

// =============================================================================
// v5: thread-per-row softmax, one warpgroup per query tile, single pass over S.
// Same CTA shape and TMEM map as v1 (two 128-row query tiles, 384 threads, S0 S1 O0 O1), but
//  * a softmax thread owns one whole query row: the 128 S values are pulled out of TMEM with four
//    tcgen05.ld in flight and ONE wait (v1/v2 made two passes with nine serialised loads, v3 splits a
//    row over two threads and pays a 256-thread barrier per tile), the row max is a FMNMX3 tree over
//    registers, exp2 runs in place and P is packed over the consumed S registers;
//  * P is handed to the tensor core in two halves (keys [0,64) and [64,128), one mbarrier each), so
//    the first four PV MMAs of a tile run while the softmax thread is still on the second half;
//  * the two warpgroups work on DIFFERENT tiles and drift half a period apart (S0 and S1 complete
//    1024 tensor cycles apart), so on every SM sub-partition one warp's TMEM load / max / TMEM store
//    phases sit under the other warp's MUFU stream.
// Why (measured, profiles/r02_k2_vs_sdpa.md): the library attention torch picks on this GPU runs
// the (T+P)^2 part at 1386 TFLOP/s where v3 reaches 1140 — v3's period per key step is twice its
// per-tile softmax time (the tiles alternate on the same 256 threads), here it is one softmax
// latency plus one QK+PV.
// =============================================================================
struct SmemV5 {
  uint8_t q[2][kTileBytes];
  uint8_t kv[kKvStages][kTileBytes];
  uint64_t q_full[2], kv_full[kKvStages], kv_empty[kKvStages], s_full[2], p_full[2][2], o_done;
  uint32_t tmem_base;
};

__device__ __forceinline__ void v5_producer(SmemV5& sm, const CUtensorMap* tmq, const CUtensorMap* tmk,
                                            const CUtensorMap* tmv, const UnitV1& u, int bh) {
  int item = 0;
  auto load_tile = [&](uint8_t* dst, const CUtensorMap* tm, uint64_t* bar, int row0) {
    mbar_expect_tx(bar, kTileBytes);
    tma_load_3d(dst, tm, bar, 0, row0, bh);
    tma_load_3d(dst + kHalfBytes, tm, bar, 64, row0, bh);
  };
  auto load_kv = [&](const CUtensorMap* tm, int j) {
    int row0, valid; key_tile(u.ks, u.n0, u.j_begin + j, row0, valid);
    const int slot = item % kKvStages;
    mbar_wait(&sm.kv_empty[slot], ((item / kKvStages) & 1) ^ 1);
    load_tile(sm.kv[slot], tm, &sm.kv_full[slot], row0);
    ++item;
  };
  load_tile(sm.q[0], tmq, &sm.q_full[0], u.q_row0[0]);
  load_kv(tmk, 0);
  if (u.n_q == 2) load_tile(sm.q[1], tmq, &sm.q_full[1], u.q_row0[1]);
  load_kv(tmv, 0);
  for (int j = 1; j < u.n_local; ++j) { load_kv(tmk, j); load_kv(tmv, j); }
}

__device__ __forceinline__ void v5_mma(SmemV5& sm, uint32_t tmem, const UnitV1& u) {
  constexpr uint32_t idesc_qk = make_idesc_bf16(kBM, kBN, false, false);
  constexpr uint32_t idesc_pv = make_idesc_bf16(kBM, kD, false, true);
  // Descriptors differ from tile to tile only in the 14-bit start-address field (bytes >> 4), so the
  // issue loop is one 32-bit add per operand: base descriptor of buffer 0 + slot * (tile bytes >> 4)
  // + the k-step offset (a compile-time constant once the loops are unrolled).
  const uint64_t dq = make_smem_desc(smem_u32(sm.q[0]), 16, 1024);
  const uint64_t dk = make_smem_desc(smem_u32(sm.kv[0]), 16, 1024);             // K tile: K-major
  const uint64_t dv = make_smem_desc(smem_u32(sm.kv[0]), kHalfBytes, 1024);     // V tile: MN-major
  auto kv_wait = [&](int item) {
    const int slot = item % kKvStages;
    mbar_wait(&sm.kv_full[slot], (item / kKvStages) & 1);
    return slot;
  };
  auto qk = [&](int t, int kslot) {                      // S_t = Q_t K^T
    const uint64_t qa = dq + (uint64_t)(t * (kTileBytes >> 4)), ka = dk + (uint64_t)(kslot * (kTileBytes >> 4));
#pragma unroll
    for (int kk = 0; kk < kD / 16; ++kk) {
      const uint32_t off = ((kk >> 2) * kHalfBytes + (kk & 3) * 32) >> 4;
      mma_ss(tmem + t * 128, qa + off, ka + off, idesc_qk, kk > 0);
    }
    mma_commit(&sm.s_full[t]);
  };
  auto pv = [&](int t, int vslot, int j) {               // O_t (+)= P_t V in two key halves
    const uint64_t va = dv + (uint64_t)(vslot * (kTileBytes >> 4));
#pragma unroll
    for (int hf = 0; hf < 2; ++hf) {
      mbar_wait(&sm.p_full[t][hf], j & 1);
      tc_fence_after();
#pragma unroll
      for (int kk = 4 * hf; kk < 4 * hf + 4; ++kk)
        mma_ts(tmem + 256 + t * 128, tmem + t * 128 + kk * 8, va + (uint64_t)(kk * (2048 >> 4)), idesc_pv,
               j > 0 || kk > 0);
    }
  };
  const bool two = u.n_q == 2;
  mbar_wait(&sm.q_full[0], 0);
  int ks = kv_wait(0);
  tc_fence_after();
  qk(0, ks);
  if (two) { mbar_wait(&sm.q_full[1], 0); tc_fence_after(); qk(1, ks); }
  mma_commit(&sm.kv_empty[ks]);
  for (int j = 0; j < u.n_local; ++j) {
    const bool has_next = j + 1 < u.n_local;
    const int vs = kv_wait(2 * j + 1);
    if (has_next) ks = kv_wait(2 * j + 2);
    tc_fence_after();
    pv(0, vs, j);
    if (has_next) qk(0, ks);
    if (two) pv(1, vs, j);
    mma_commit(&sm.kv_empty[vs]);
    if (has_next) {
      if (two) qk(1, ks);
      mma_commit(&sm.kv_empty[ks]);
    }
  }
  mma_commit(&sm.o_done);
}

template <int kPolyEvery>
__device__ __forceinline__ void v5_softmax(SmemV5& sm, uint32_t tmem, const UnitV1& u,
                                           const AttnParams& p, int t, int b, int h) {
  const int warp_q = (threadIdx.x >> 5) & 3, lane = threadIdx.x & 31;
  const int row = warp_q * 32 + lane;
  const uint32_t lane_off = (uint32_t)(warp_q * 32) << 16;
  const uint32_t t_s = tmem + t * 128 + lane_off, t_o = tmem + 256 + t * 128 + lane_off;
  const float scale = p.scale_log2;
  const int q_row0 = t ? u.q_row0[1] : u.q_row0[0], q_rows = t ? u.q_rows[1] : u.q_rows[0];
  float m_used = -INFINITY, l_sum = 0.f;

#pragma unroll 1
  for (int j = 0; j < u.n_local; ++j) {
    int row0, valid; key_tile(u.ks, u.n0, u.j_begin + j, row0, valid);
    mbar_wait(&sm.s_full[t], j & 1);
    tc_fence_after();
    uint32_t s[kBN];
    tmem_ld32(t_s, s);
    tmem_ld32(t_s + 32, s + 32);
    tmem_ld32(t_s + 64, s + 64);
    tmem_ld32(t_s + 96, s + 96);
    tmem_wait_ld32(s);
    tmem_wait_ld32(s + 32);
    tmem_wait_ld32(s + 64);
    tmem_wait_ld32(s + 96);
    if (valid < kBN) {                                      // ragged last tile of a key segment
#pragma unroll
      for (int i = 0; i < kBN; ++i) if (i >= valid) s[i] = 0xff800000u;   // -inf
    }
    // row max: four independent FMNMX3 chains
    float mc[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) mc[c] = __uint_as_float(s[c]);
#pragma unroll
    for (int i = 4; i < kBN; i += 8) {
#pragma unroll
      for (int c = 0; c < 4; ++c)
        mc[c] = fmaxf(fmaxf(mc[c], __uint_as_float(s[i + c])), __uint_as_float(s[(i + 4 + c) & (kBN - 1)]));
    }
    const float mx = fmaxf(fmaxf(mc[0], mc[1]), fmaxf(mc[2], mc[3]));
    const bool grow = (mx - m_used) * scale > kRescaleThreshold;   // true on the first tile
    const float m_new = grow ? mx : m_used;
    if (j > 0 && __any_sync(0xffffffffu, grow)) {
      // PV_t(j-1) has completed: the commit behind s_full[t] covers it.  Rare (lazy threshold).
      const float alpha = grow ? ex2((m_used - m_new) * scale) : 1.f;
      l_sum *= alpha;
#pragma unroll 1
      for (int c = 0; c < kD; c += 8) {
        uint32_t rc[8];
        tmem_ld8(t_o + c, rc);
        tmem_wait_ld8(rc);
#pragma unroll
        for (int i = 0; i < 8; ++i) rc[i] = __float_as_uint(__uint_as_float(rc[i]) * alpha);
        tmem_st8(t_o + c, rc);
      }
      tmem_wait_st();
    }
    m_used = m_new;
    const float mb = m_used * scale;
    const uint64_t scale2 = pk64(scale, scale), nmb2 = pk64(-mb, -mb);
    uint64_t rsa = pk64(0.f, 0.f), rsb = pk64(0.f, 0.f);
#pragma unroll
    for (int hf = 0; hf < 2; ++hf) {
      uint32_t pk[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const int e = 64 * hf + 2 * i;
        const uint64_t x2 = fma2(pk64(__uint_as_float(s[e]), __uint_as_float(s[e + 1])), scale2, nmb2);
        float e0, e1;
        if (use_poly<kPolyEvery>(i)) {
          exp2_poly2(x2, e0, e1);
        } else {
          float x0, x1;
          upk64(x2, x0, x1);
          e0 = ex2(x0); e1 = ex2(x1);
        }
        if (i & 1) rsb = add2(rsb, pk64(e0, e1)); else rsa = add2(rsa, pk64(e0, e1));
        pk[i] = pack2(e0, e1);
      }
      tmem_st32(t_s + hf * 32, pk);                       // P_t keys [64 hf, 64 hf + 64) over S_t columns
      tmem_wait_st();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm.p_full[t][hf]);
    }
    {
      float r0, r1, r2, r3;
      upk64(rsa, r0, r1); upk64(rsb, r2, r3);
      l_sum += (r0 + r1) + (r2 + r3);
    }
  }

  // ---- epilogue: O_t / l -> bf16 "B L (H D)" ---------------------------------------
  mbar_wait(&sm.o_done, 0);
  tc_fence_after();
  const float inv_l = 1.f / l_sum;
#pragma unroll
  for (int c = 0; c < kD; c += 32) {
    uint32_t r[32];
    tmem_ld32(t_o + c, r);
    tmem_wait_ld32(r);
    if (row < q_rows) {
      float o[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) o[i] = __uint_as_float(r[i]);
      store_out_row(p, b, h, q_row0 + row, o, inv_l, c, 32);
    }
  }
}

template <int kPolyEvery>
__global__ void __launch_bounds__(384, 1)
attn_v5_kernel(const __grid_constant__ CUtensorMap tmq, const __grid_constant__ CUtensorMap tmk,
               const __grid_constant__ CUtensorMap tmv, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  SmemV5& sm = *reinterpret_cast<SmemV5*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const UnitV1 u = decode_unit(p, blockIdx.x);
  if (u.n_local == 0) return;
  const int bh = u.bh, b = bh / p.H, h = bh - b * p.H;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmq); prefetch_tmap(&tmk); prefetch_tmap(&tmv);
    mbar_init(&sm.q_full[0], 1); mbar_init(&sm.q_full[1], 1);
    for (int i = 0; i < kKvStages; ++i) { mbar_init(&sm.kv_full[i], 1); mbar_init(&sm.kv_empty[i], 1); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&sm.s_full[i], 1);
      mbar_init(&sm.p_full[i][0], 4); mbar_init(&sm.p_full[i][1], 4);
    }
    mbar_init(&sm.o_done, 1);
    fence_barrier_init();
  }
  if (warp == 2) { tmem_alloc(&sm.tmem_base, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  if (warp < 4) {
    // register pool of a 384-thread CTA: 384 x 168 = 64512 = 128 x 88 + 256 x 208
    setmaxnreg_dec<88>();
    if (warp == 0) {
      if (elect_one()) v5_producer(sm, &tmq, &tmk, &tmv, u, bh);
    } else if (warp == 1) {
      if (elect_one()) v5_mma(sm, tmem, u);
    }
  } else {
    setmaxnreg_inc<208>();
    const int t = (warp >> 2) - 1;
    if (t < u.n_q) v5_softmax<kPolyEvery>(sm, tmem, u, p, t, b, h);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem, 512);
}

// =============================================================================
// v6: v5 with S double-buffered per query tile.  The 128 TMEM columns of S_t hold TWO 64-key
// score buffers; a key tile is processed as two chunks c = 2j + h:
//     S_t[h] = Q_t K[64h:64h+64]^T   (8 x tcgen05.mma N = 64)
//     O_t   += P_t[h] V[64h:64h+64]  (4 x tcgen05.mma, P over the first 32 columns of S_t[h])
// and the MMA warp runs ONE chunk ahead of the softmax:  QK(c+2) is issued right behind PV(c), so
// while a softmax thread works on chunk c the scores of chunk c+1 are already in TMEM (v5: the
// warpgroup idles through PV + QK of its own tile, 36 % of its samples on the s_full wait).  Both
// warpgroups stay busy, the tensor pipe always has a queued chunk of the other buffer.
// The one ordering this breaks: s_full of chunk c no longer implies that PV(c-1) has completed,
// so the (rare, lazy) rescale of O waits on pv_done[t], committed behind every PV.
// =============================================================================
struct SmemV6 {
  uint8_t q[2][kTileBytes];
  uint8_t kv[kKvStages][kTileBytes];
  uint64_t q_full[2], kv_full[kKvStages], kv_empty[kKvStages], s_full[2][2], p_full[2][2], pv_done[2], o_done;
  uint32_t tmem_base;
};

__device__ __forceinline__ void v6_mma(SmemV6& sm, uint32_t tmem, const UnitV1& u) {
  constexpr uint32_t idesc_qk = make_idesc_bf16(kBM, 64, false, false);
  constexpr uint32_t idesc_pv = make_idesc_bf16(kBM, kD, false, true);
  const uint64_t dq = make_smem_desc(smem_u32(sm.q[0]), 16, 1024);
  const uint64_t dk = make_smem_desc(smem_u32(sm.kv[0]), 16, 1024);             // K tile: K-major
  const uint64_t dv = make_smem_desc(smem_u32(sm.kv[0]), kHalfBytes, 1024);     // V tile: MN-major
  auto kv_wait = [&](int item) {
    const int slot = item % kKvStages;
    mbar_wait(&sm.kv_full[slot], (item / kKvStages) & 1);
    return slot;
  };
  auto qk = [&](int t, int kslot, int hf) {              // S_t[hf] = Q_t K[64 hf : 64 hf + 64]^T
    const uint64_t qa = dq + (uint64_t)(t * (kTileBytes >> 4));
    const uint64_t ka = dk + (uint64_t)(kslot * (kTileBytes >> 4) + hf * ((64 * 128) >> 4));
#pragma unroll
    for (int kk = 0; kk < kD / 16; ++kk) {
      const uint32_t off = ((kk >> 2) * kHalfBytes + (kk & 3) * 32) >> 4;
      mma_ss(tmem + t * 128 + hf * 64, qa + off, ka + off, idesc_qk, kk > 0);
    }
    mma_commit(&sm.s_full[t][hf]);
  };
  auto pv = [&](int t, int vslot, int j, int hf) {       // O_t (+)= P_t[hf] V[64 hf : 64 hf + 64]
    const uint64_t va = dv + (uint64_t)(vslot * (kTileBytes >> 4) + hf * ((4 * 2048) >> 4));
    mbar_wait(&sm.p_full[t][hf], j & 1);
    tc_fence_after();
#pragma unroll
    for (int kk = 0; kk < 4; ++kk)
      mma_ts(tmem + 256 + t * 128, tmem + t * 128 + hf * 64 + kk * 8, va + (uint64_t)(kk * (2048 >> 4)), idesc_pv,
             j > 0 || hf > 0 || kk > 0);
    mma_commit(&sm.pv_done[t]);
  };
  const bool two = u.n_q == 2;
  mbar_wait(&sm.q_full[0], 0);
  int ks = kv_wait(0);
  tc_fence_after();
  qk(0, ks, 0);
  if (two) { mbar_wait(&sm.q_full[1], 0); tc_fence_after(); qk(1, ks, 0); }
  qk(0, ks, 1);
  if (two) qk(1, ks, 1);
  mma_commit(&sm.kv_empty[ks]);
  for (int j = 0; j < u.n_local; ++j) {
    const bool has_next = j + 1 < u.n_local;
    const int vs = kv_wait(2 * j + 1);
    if (has_next) ks = kv_wait(2 * j + 2);
    tc_fence_after();
#pragma unroll
    for (int hf = 0; hf < 2; ++hf) {
      pv(0, vs, j, hf);
      if (has_next) qk(0, ks, hf);
      if (two) {
        pv(1, vs, j, hf);
        if (has_next) qk(1, ks, hf);
      }
    }
    mma_commit(&sm.kv_empty[vs]);
    if (has_next) mma_commit(&sm.kv_empty[ks]);
  }
  mma_commit(&sm.o_done);
}

__device__ __forceinline__ void v6_producer(SmemV6& sm, const CUtensorMap* tmq, const CUtensorMap* tmk,
                                            const CUtensorMap* tmv, const UnitV1& u, int bh) {
  int item = 0;
  auto load_tile = [&](uint8_t* dst, const CUtensorMap* tm, uint64_t* bar, int row0) {
    mbar_expect_tx(bar, kTileBytes);
    tma_load_3d(dst, tm, bar, 0, row0, bh);
    tma_load_3d(dst + kHalfBytes, tm, bar, 64, row0, bh);
  };
  auto load_kv = [&](const CUtensorMap* tm, int j) {
    int row0, valid; key_tile(u.ks, u.n0, u.j_begin + j, row0, valid);
    const int slot = item % kKvStages;
    mbar_wait(&sm.kv_empty[slot], ((item / kKvStages) & 1) ^ 1);
    load_tile(sm.kv[slot], tm, &sm.kv_full[slot], row0);
    ++item;
  };
  load_tile(sm.q[0], tmq, &sm.q_full[0], u.q_row0[0]);
  load_kv(tmk, 0);
  if (u.n_q == 2) load_tile(sm.q[1], tmq, &sm.q_full[1], u.q_row0[1]);
  load_kv(tmv, 0);
  for (int j = 1; j < u.n_local; ++j) { load_kv(tmk, j); load_kv(tmv, j); }
}

template <int kPolyEvery>
__device__ __forceinline__ void v6_softmax(SmemV6& sm, uint32_t tmem, const UnitV1& u,
                                           const AttnParams& p, int t, int b, int h) {
  const int warp_q = (threadIdx.x >> 5) & 3, lane = threadIdx.x & 31;
  const int row = warp_q * 32 + lane;
  const uint32_t lane_off = (uint32_t)(warp_q * 32) << 16;
  const uint32_t t_s = tmem + t * 128 + lane_off, t_o = tmem + 256 + t * 128 + lane_off;
  const float scale = p.scale_log2;
  const int q_row0 = t ? u.q_row0[1] : u.q_row0[0], q_rows = t ? u.q_rows[1] : u.q_rows[0];
  float m_used = -INFINITY, l_sum = 0.f;
  const uint64_t scale2 = pk64(scale, scale);

#pragma unroll 1
  for (int j = 0; j < u.n_local; ++j) {
    int row0, valid; key_tile(u.ks, u.n0, u.j_begin + j, row0, valid);
#pragma unroll
    for (int hf = 0; hf < 2; ++hf) {
      mbar_wait(&sm.s_full[t][hf], j & 1);
      tc_fence_after();
      uint32_t s[64];
      tmem_ld32(t_s + hf * 64, s);
      tmem_ld32(t_s + hf * 64 + 32, s + 32);
      tmem_wait_ld32(s);
      tmem_wait_ld32(s + 32);
      if (valid < kBN) {                                    // ragged last tile of a key segment
#pragma unroll
        for (int i = 0; i < 64; ++i) if (hf * 64 + i >= valid) s[i] = 0xff800000u;   // -inf
      }
      float mc[4];
#pragma unroll
      for (int c = 0; c < 4; ++c) mc[c] = __uint_as_float(s[c]);
#pragma unroll
      for (int i = 4; i < 64; i += 8) {
#pragma unroll
        for (int c = 0; c < 4; ++c)
          mc[c] = fmaxf(fmaxf(mc[c], __uint_as_float(s[i + c])), __uint_as_float(s[(i + 4 + c) & 63]));
      }
      const float mx = fmaxf(fmaxf(mc[0], mc[1]), fmaxf(mc[2], mc[3]));
      // a chunk that lies wholly past the segment end (valid <= 64, hf = 1) has mx = -inf: nothing grows
      const bool grow = (mx - m_used) * scale > kRescaleThreshold;   // true on the first chunk
      const float m_new = grow ? mx : m_used;
      const int c_idx = 2 * j + hf;
      if (c_idx > 0 && __any_sync(0xffffffffu, grow)) {
        // rare (lazy threshold): O must hold every PV up to chunk c-1 before it is rescaled
        mbar_wait(&sm.pv_done[t], (c_idx - 1) & 1);
        tc_fence_after();
        const float alpha = grow ? ex2((m_used - m_new) * scale) : 1.f;
        l_sum *= alpha;
#pragma unroll 1
        for (int c = 0; c < kD; c += 8) {
          uint32_t rc[8];
          tmem_ld8(t_o + c, rc);
          tmem_wait_ld8(rc);
#pragma unroll
          for (int i = 0; i < 8; ++i) rc[i] = __float_as_uint(__uint_as_float(rc[i]) * alpha);
          tmem_st8(t_o + c, rc);
        }
        tmem_wait_st();
      }
      m_used = m_new;
      const float mb = m_used * scale;
      const uint64_t nmb2 = pk64(-mb, -mb);
      uint64_t rsa = pk64(0.f, 0.f), rsb = pk64(0.f, 0.f);
      uint32_t pk[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const uint64_t x2 = fma2(pk64(__uint_as_float(s[2 * i]), __uint_as_float(s[2 * i + 1])), scale2, nmb2);
        float e0, e1;
        if (use_poly<kPolyEvery>(i)) {
          exp2_poly2(x2, e0, e1);
        } else {
          float x0, x1;
          upk64(x2, x0, x1);
          e0 = ex2(x0); e1 = ex2(x1);
        }
        if (i & 1) rsb = add2(rsb, pk64(e0, e1)); else rsa = add2(rsa, pk64(e0, e1));
        pk[i] = pack2(e0, e1);
      }
      tmem_st32(t_s + hf * 64, pk);                       // P_t[hf] over the first 32 columns of S_t[hf]
      tmem_wait_st();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm.p_full[t][hf]);
      float r0, r1, r2, r3;
      upk64(rsa, r0, r1); upk64(rsb, r2, r3);
      l_sum += (r0 + r1) + (r2 + r3);
    }
  }

  // ---- epilogue: O_t / l -> bf16 "B L (H D)" ---------------------------------------
  mbar_wait(&sm.o_done, 0);
  tc_fence_after();
  const float inv_l = 1.f / l_sum;
#pragma unroll
  for (int c = 0; c < kD; c += 32) {
    uint32_t r[32];
    tmem_ld32(t_o + c, r);
    tmem_wait_ld32(r);
    if (row < q_rows) {
      float o[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) o[i] = __uint_as_float(r[i]);
      store_out_row(p, b, h, q_row0 + row, o, inv_l, c, 32);
    }
  }
}

template <int kPolyEvery>
__global__ void __launch_bounds__(384, 1)
attn_v6_kernel(const __grid_constant__ CUtensorMap tmq, const __grid_constant__ CUtensorMap tmk,
               const __grid_constant__ CUtensorMap tmv, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  SmemV6& sm = *reinterpret_cast<SmemV6*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const UnitV1 u = decode_unit(p, blockIdx.x);
  if (u.n_local == 0) return;
  const int bh = u.bh, b = bh / p.H, h = bh - b * p.H;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmq); prefetch_tmap(&tmk); prefetch_tmap(&tmv);
    mbar_init(&sm.q_full[0], 1); mbar_init(&sm.q_full[1], 1);
    for (int i = 0; i < kKvStages; ++i) { mbar_init(&sm.kv_full[i], 1); mbar_init(&sm.kv_empty[i], 1); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&sm.s_full[i][0], 1); mbar_init(&sm.s_full[i][1], 1);
      mbar_init(&sm.p_full[i][0], 4); mbar_init(&sm.p_full[i][1], 4);
      mbar_init(&sm.pv_done[i], 1);
    }
    mbar_init(&sm.o_done, 1);
    fence_barrier_init();
  }
  if (warp == 2) { tmem_alloc(&sm.tmem_base, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  if (warp < 4) {
    // register pool of a 384-thread CTA: 384 x 168 = 64512 >= 128 x 96 + 256 x 200
    setmaxnreg_dec<96>();
    if (warp == 0) {
      if (elect_one()) v6_producer(sm, &tmq, &tmk, &tmv, u, bh);
    } else if (warp == 1) {
      if (elect_one()) v6_mma(sm, tmem, u);
    }
  } else {
    setmaxnreg_inc<200>();
    const int t = (warp >> 2) - 1;
    if (t < u.n_q) v6_softmax<kPolyEvery>(sm, tmem, u, p, t, b, h);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem, 512);
}

#include "k2_attention_v7.cuh"

// CA_ATTN_NEITHER: concept_attn = concept_v  (modified_double_stream_block.py:157-159),
// i.e. a head-merge copy of the concept rows of V.
__global__ void concept_v_copy_kernel(const __nv_bfloat16* __restrict__ v, AttnParams p) {
  const int Ntot = p.P + p.T + p.C;
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;   // one 16-byte chunk each
  const int64_t total = (int64_t)p.B * p.H * p.C * (kD / 8);
  if (i >= total) return;
  const int ch = (int)(i % (kD / 8));
  const int c = (int)((i / (kD / 8)) % p.C);
  const int h = (int)((i / ((kD / 8) * p.C)) % p.H);
  const int b = (int)(i / ((int64_t)(kD / 8) * p.C * p.H));
  const uint4 val = *reinterpret_cast<const uint4*>(
      v + (((int64_t)b * p.H + h) * Ntot + p.P + p.T + c) * kD + ch * 8);
  *reinterpret_cast<uint4*>(p.out_txtc + ((int64_t)b * (p.T + p.C) + p.T + c) * p.H * kD + h * kD +
                            ch * 8) = val;
}

// ---- host side -------------------------------------------------------------------
// [B*H, Ntot, 128] bf16 viewed as a 3-D tensor; box = 64 x box_rows x 1, 128B swizzle.
static ca_status make_qkv_tmap(CUtensorMap* m, const void* base, int BH, int Ntot, int box_rows) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) return CA_ERR_CUDA;
  const cuuint64_t dims[3] = {(cuuint64_t)kD, (cuuint64_t)Ntot, (cuuint64_t)BH};
  const cuuint64_t strides[2] = {(cuuint64_t)kD * 2, (cuuint64_t)Ntot * kD * 2};
  const cuuint32_t box[3] = {64, (cuuint32_t)box_rows, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides,
                   box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { g_last_cuda_error = (int)r; return CA_ERR_CUDA; }
  return CA_OK;
}

int g_attention_impl = 0;    // 0 = best available, 1 = force v0 (debug / A-B tests)
int g_attention_streamk = 0; // 0 = auto (real SM count), -1 = never split, n > 0 = plan for n SMs (tests)

// Stream-K plan for the main units of one launch.  W units of n key steps each on S SMs: the
// first floor(W/S)*S run whole (one CTA each); the remaining R < S would leave S-R SMs idle for a
// whole unit time, so their R*n key steps are cut into S equal ranges instead (each range touches
// at most two units -> two CTAs per range, pieces of a unit merged by the last one to finish).
struct SkPlan { int n_dp, sk_units, sk_sms; };

static SkPlan plan_streamk_core(int W, int n_concept, int n_steps, int S, bool auto_rules);

static SkPlan plan_streamk(int W, int n_concept, int n_steps, bool allowed) {
  SkPlan pl{W, 0, 0};
  if (!allowed || g_attention_streamk < 0 || W == 0) return pl;
  int S = g_attention_streamk;
  if (S == 0) {
    static int sm_count = 0;
    if (sm_count == 0) {
      int dev = 0;
      if (cudaGetDevice(&dev) != cudaSuccess ||
          cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
        return pl;
    }
    S = sm_count;
  }
  return plan_streamk_core(W, n_concept, n_steps, S, g_attention_streamk == 0);
}

// pure arithmetic (no device, no globals): S = SMs planned for, auto_rules = the measured thresholds
static SkPlan plan_streamk_core(int W, int n_concept, int n_steps, int S, bool auto_rules) {
  SkPlan pl{W, 0, 0};
  if (W <= 0 || S <= 0) return pl;
  // the concept units (one CTA each, launched first) count towards the whole waves
  const int whole = ((W + n_concept) / S) * S - n_concept;      // main units that run unsplit
  const int R = whole >= 0 ? W - whole : W;
  if (R <= 0 || R >= S) return pl;
  // Splitting saves up to (1 - R/S) * n_steps key steps of wall time on the last wave.  It costs
  // two extra CTA start-ups per SM, a 133 KB partial per piece and the merge (~10 key steps), and
  // the two pieces of a range are separate CTAs placed by the hardware scheduler, so the last wave
  // ends after about 1.5 range lengths, not 1.  Measured on B200 (scripts/bench_attention.py):
  // R/S = 0.34 at 130 steps: 0.488 -> 0.394 ms; R/S = 0.35 (12 heads): 1.420 -> 1.391 ms;
  // R/S = 0.70 at 130 steps: 2.780 -> 2.856 ms (worse); R/S = 0.76 at 34 steps: worse.
  // Hence: split only when the tail wave is at most half full and the saving is >= 24 steps.
  // In test mode (planned SM count given) only the structural limits apply.
  const int64_t saved = (int64_t)(S - R) * n_steps / S;
  if ((int64_t)R * n_steps / S < 4) return pl;
  if (auto_rules && (2 * R > S || saved < 24)) return pl;
  pl.n_dp = W - R; pl.sk_units = R; pl.sk_sms = S;
  return pl;
}

static bool impl_is_v3() {
  return g_attention_impl == 0 || g_attention_impl == 4 || g_attention_impl == 5 || g_attention_impl == 10 ||
         g_attention_impl == 13 || g_attention_impl == 14;
}

static int sm_count_cached() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
      n = 0;
  }
  return n;
}

}  // namespace ca

extern "C" void ca_debug_set_attention_impl(int impl) { ca::g_attention_impl = impl; }
extern "C" void ca_debug_set_attention_streamk(int sms) { ca::g_attention_streamk = sms; }

// Host-only views of the stream-K arithmetic for the CPU tests (no CUDA call inside).
// plan: out = {n_dp, sk_units, sk_sms};  piece: out = {unit, j_begin, n_local}, returns 0 for an empty piece.
extern "C" void ca_debug_plan_streamk(int W, int n_concept, int n_steps, int sms, int auto_rules, int* out) {
  const ca::SkPlan pl = ca::plan_streamk_core(W, n_concept, n_steps, sms, auto_rules != 0);
  out[0] = pl.n_dp; out[1] = pl.sk_units; out[2] = pl.sk_sms;
}
extern "C" int ca_debug_streamk_piece(int n_steps, int sk_units, int sk_sms, int i, int k, int* out) {
  ca::AttnParams p{};
  p.sk_units = sk_units; p.sk_sms = sk_sms;
  int unit = 0, j_begin = 0, n_local = 0;
  const bool ok = ca::sk_piece(p, n_steps, i, k, unit, j_begin, n_local);
  out[0] = unit; out[1] = j_begin; out[2] = n_local;
  return ok ? 1 : 0;
}

extern "C" size_t ca_joint_attention_workspace_bytes(int B, int H, int P, int T, int C, int mode) {
  using namespace ca;
  if (B <= 0 || H <= 0 || P + T <= 0 || !impl_is_v3()) return 0;
  const int n_main = (P + T + kBM - 1) / kBM;
  const int n_concept = (C > 0 && mode != CA_ATTN_NEITHER) ? B * H : 0;
  const SkPlan pl = plan_streamk(((n_main + 1) / 2) * B * H, n_concept, n_main, true);
  if (pl.sk_sms == 0) return 0;
  return (size_t)2 * pl.sk_sms * kPartFloats * sizeof(float) + (size_t)pl.sk_units * sizeof(int) + 256;
}

extern "C" ca_status ca_joint_attention_ws(const void* q, const void* k, const void* v, int B, int H,
                                           int P, int T, int C, int mode, void* out_img,
                                           void* out_txtc, void* workspace, size_t workspace_bytes,
                                           void* stream) {
  using namespace ca;
  if (!q || !k || !v || B < 0 || H <= 0 || P < 0 || T < 0 || C < 0) return CA_ERR_INVALID_ARG;
  if (mode < CA_ATTN_BOTH || mode > CA_ATTN_NEITHER) return CA_ERR_INVALID_ARG;
  if (C > CA_MAX_CONCEPTS) return CA_ERR_UNSUPPORTED;
  if ((P > 0 && !out_img) || (T + C > 0 && !out_txtc)) return CA_ERR_INVALID_ARG;
  const void* ptrs[] = {q, k, v, out_img, out_txtc, workspace};
  for (const void* ptr : ptrs) if (ptr && !aligned16(ptr)) return CA_ERR_ALIGNMENT;
  const int Ntot = P + T + C;
  if (B == 0 || Ntot == 0) return CA_OK;
  if ((int64_t)B * H > 65535) return CA_ERR_UNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;

  AttnParams p;
  p.B = B; p.H = H; p.P = P; p.T = T; p.C = C; p.mode = mode;
  p.n_main_tiles = (P + T + kBM - 1) / kBM;
  p.out_img = (__nv_bfloat16*)out_img;
  p.out_txtc = (__nv_bfloat16*)out_txtc;
  p.scale_log2 = kLog2e / sqrtf((float)kD);
  p.n_pairs = (p.n_main_tiles + 1) / 2;
  p.ws_part = nullptr; p.ws_count = nullptr;

  CUtensorMap tmq, tmk, tmv;
  ca_status s;
  if ((s = make_qkv_tmap(&tmq, q, B * H, Ntot, kBM)) != CA_OK) return s;
  if ((s = make_qkv_tmap(&tmk, k, B * H, Ntot, kBN)) != CA_OK) return s;
  if ((s = make_qkv_tmap(&tmv, v, B * H, Ntot, kBN)) != CA_OK) return s;

  const bool concept_attn = C > 0 && mode != CA_ATTN_NEITHER;
  if (g_attention_impl == 1) {
    const int units = p.n_main_tiles + (concept_attn ? 1 : 0);
    if (units > 0) {
      const int smem = (int)sizeof(SmemV0) + 1024;
      CA_CUDA_OK(cudaFuncSetAttribute(attn_v0_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      attn_v0_kernel<<<dim3(units, B * H), 128, smem, st>>>(tmq, tmk, tmv, p);
      CA_LAUNCH_CHECK();
    }
  } else {
    const int W = p.n_pairs * B * H;
    const size_t need = ca_joint_attention_workspace_bytes(B, H, P, T, C, mode);
    p.n_concept = concept_attn ? B * H : 0;
    const SkPlan pl = plan_streamk(W, p.n_concept, p.n_main_tiles,
                                   impl_is_v3() && workspace != nullptr && need > 0 && workspace_bytes >= need);
    p.n_dp = pl.n_dp; p.sk_units = pl.sk_units; p.sk_sms = pl.sk_sms;
    if (pl.sk_sms > 0) {
      p.ws_part = (float*)workspace;
      p.ws_count = (int*)((char*)workspace + (size_t)2 * pl.sk_sms * kPartFloats * sizeof(float));
      CA_CUDA_OK(cudaMemsetAsync(p.ws_count, 0, (size_t)pl.sk_units * sizeof(int), st));
    }
    const int64_t ctas = (int64_t)p.n_dp + 2 * p.sk_sms + p.n_concept;
    if (ctas > 0x7fffffffLL) return CA_ERR_UNSUPPORTED;
    if (ctas > 0 && g_attention_impl >= 50 && g_attention_impl <= 52 && pl.sk_sms == 0) {
      // v11 (one score buffer shared by the two tiles, P in TMEM, five K/V stages): 50 / 51 / 52 as 40 / 41 / 42
      const int smem = (int)sizeof(SmemV11) + 1024;
      auto kern = g_attention_impl == 51 ? attn_v11_kernel<0> : g_attention_impl == 52 ? attn_v11_kernel<2>
                                                              : attn_v11_kernel<4>;
      CA_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      kern<<<(unsigned)ctas, 384, smem, st>>>(tmq, tmk, tmv, p);
      CA_LAUNCH_CHECK();
    } else if (ctas > 0 && g_attention_impl >= 40 && g_attention_impl <= 45 && pl.sk_sms == 0) {
      // v7 (P through shared memory, S freed when loaded): 40 = every 4th pair on the polynomial,
      // 41 = all MUFU, 42 = every 2nd;  v10 = v7 with the MMA issue split over three threads: 43 / 44 / 45
      const int smem = (int)sizeof(SmemV7) + 1024;
      auto kern = g_attention_impl == 40 ? attn_v7_kernel<4, false> : g_attention_impl == 41 ? attn_v7_kernel<0, false>
                : g_attention_impl == 42 ? attn_v7_kernel<2, false> : g_attention_impl == 43 ? attn_v7_kernel<4, true>
                : g_attention_impl == 44 ? attn_v7_kernel<0, true> : attn_v7_kernel<2, true>;
      CA_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      kern<<<(unsigned)ctas, 384, smem, st>>>(tmq, tmk, tmv, p);
      CA_LAUNCH_CHECK();
    } else if (ctas > 0 && g_attention_impl >= 30 && g_attention_impl <= 32 && pl.sk_sms == 0) {
      // v6 (thread-per-row softmax, S double-buffered in 64-key chunks): 30 = every 4th pair on the
      // polynomial, 31 = all MUFU, 32 = every 2nd
      const int smem = (int)sizeof(SmemV6) + 1024;
      auto kern = g_attention_impl == 31 ? attn_v6_kernel<0> : g_attention_impl == 32 ? attn_v6_kernel<2>
                                                             : attn_v6_kernel<4>;
      CA_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      kern<<<(unsigned)ctas, 384, smem, st>>>(tmq, tmk, tmv, p);
      CA_LAUNCH_CHECK();
    } else if (ctas > 0 && g_attention_impl >= 20 && g_attention_impl <= 22 && pl.sk_sms == 0) {
      // v5 (thread-per-row softmax): 20 = every 4th pair on the polynomial, 21 = all MUFU, 22 = every 2nd
      const int smem = (int)sizeof(SmemV5) + 1024;
      auto kern = g_attention_impl == 21 ? attn_v5_kernel<0> : g_attention_impl == 22 ? attn_v5_kernel<2>
                                                             : attn_v5_kernel<4>;
      CA_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      kern<<<(unsigned)ctas, 384, smem, st>>>(tmq, tmk, tmv, p);
      CA_LAUNCH_CHECK();
    } else if (ctas > 0 && g_attention_impl == 14) {
      // experimental persistent v3: one CTA per SM walks the same flat unit list
      const int sms = sm_count_cached();
      if (sms <= 0) return CA_ERR_CUDA;
      const int smem = (int)sizeof(SmemP) + 1024;
      auto kern = attn_persist_kernel<4>;
      CA_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      kern<<<(unsigned)(ctas < sms ? ctas : sms), 384, smem, st>>>(tmq, tmk, tmv, p, (int)ctas);
      CA_LAUNCH_CHECK();
    } else if (ctas > 0) {
      const int smem = (int)sizeof(SmemV1) + 1024;
      // impl 0: v3 (two cooperating warpgroups), polynomial exp2 on every 4th element pair;
      // 4 / 5: v3 every 2nd pair / all MUFU; 2 / 3: v2 (one warpgroup per tile) all MUFU / every 4th;
      // 10: v3 with 3 of every 8 pairs on the polynomial;
      // 9 / 6 / 7: v4 (four warpgroups, 640 threads) every 4th / every 2nd / all MUFU — measured
      // 3-5 % slower than v3 on B200 (profiles/r01_k2_v4_experiment.md), kept for A/B runs.
      auto kern = g_attention_impl == 2 ? attn_v1_kernel<0, 0>
                : g_attention_impl == 3 ? attn_v1_kernel<4, 0>
                : g_attention_impl == 4 ? attn_v1_kernel<2, 1>
                : g_attention_impl == 5 ? attn_v1_kernel<0, 1>
                : g_attention_impl == 9 ? attn_v1_kernel<4, 2>
                : g_attention_impl == 10 ? attn_v1_kernel<-3, 1>
                : g_attention_impl == 13 ? attn_v1_kernel<4, 3>
                : g_attention_impl == 6 ? attn_v1_kernel<2, 2>
                : g_attention_impl == 7 ? attn_v1_kernel<0, 2> : attn_v1_kernel<4, 1>;
      const bool v4 = g_attention_impl == 9 || g_attention_impl == 6 || g_attention_impl == 7;
      CA_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      kern<<<(unsigned)ctas, v4 ? 640 : 384, smem, st>>>(tmq, tmk, tmv, p);
      CA_LAUNCH_CHECK();
    }
  }
  if (C > 0 && mode == CA_ATTN_NEITHER) {
    const int64_t total = (int64_t)B * H * C * (kD / 8);
    concept_v_copy_kernel<<<(unsigned)((total + 127) / 128), 128, 0, st>>>((const __nv_bfloat16*)v, p);
    CA_LAUNCH_CHECK();
  }
  return CA_OK;
}

extern "C" ca_status ca_joint_attention(const void* q, const void* k, const void* v, int B, int H,
                                        int P, int T, int C, int mode, void* out_img,
                                        void* out_txtc, void* stream) {
  return ca_joint_attention_ws(q, k, v, B, H, P, T, C, mode, out_img, out_txtc, nullptr, 0, stream);
}
