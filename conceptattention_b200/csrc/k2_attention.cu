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
  // persistent v11 only: key-step ranges instead of whole units (k2_attention_v7.cuh, "v11p stream-K")
  int skp_on;                          // 1: CTA c works on main key steps [B(c), B(c+1))
  int skp_wc;                          // cost of one concept unit in main key steps (balancing weight)
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
  float xchg[2][2][kBM];        // v3: [parity][half][row] row max / row sum exchange
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

__device__ __forceinline__ void softmax_bar_sync(int id) {
  asm volatile("bar.sync %0, 256;" ::"r"(id) : "memory");
}

// v3 softmax: the 256 softmax threads (two warpgroups) walk tile 0 then tile 1 of every key
// step; a query row is shared by two threads (keys [0,64) and [64,128)), which agree on the row
// max through shared memory and one 256-thread named barrier per tile.
template <int kPolyEvery>
__device__ __forceinline__ void v3_softmax(SmemV1& sm, uint32_t tmem, const UnitV1& u,
                                           const AttnParams& p, int half, int b, int h) {
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
    float* xb = &sm.xchg[use & 1][0][0];
    xb[half * kBM + row] = mine;
    softmax_bar_sync(1);
    const float other = xb[(half ^ 1) * kBM + row];
    ++use;
    return take_max ? fmaxf(mine, other) : mine + other;
  };

#pragma unroll 1
  for (int j = 0; j < u.n_local; ++j) {
    int row0, valid; key_tile(u.ks, u.n0, u.j_begin + j, row0, valid);
#pragma unroll
    for (int t = 0; t < 2; ++t) {
      if (t < u.n_q) {
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
    return;
  }

  // ---- stream-K piece: this CTA saw only key tiles [j_begin, j_begin + n_local) of its unit.
  // Write (O unnormalised, row max, row sum) to the piece's slot; the piece that arrives last
  // (atomic ticket, no waiting) merges all pieces of the unit in a fixed order and stores.
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
  float* xflag = &sm.xchg[0][0][0];
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

// v3 kernel (round 1's default; the only one with the stream-K tail): CA_K2_V3.
template <int kPolyEvery>
__global__ void __launch_bounds__(384, 1)
attn_v3_kernel(const __grid_constant__ CUtensorMap tmq, const __grid_constant__ CUtensorMap tmk,
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
    for (int i = 0; i < 2; ++i) { mbar_init(&sm.s_full[i], 1); mbar_init(&sm.p_full[i], 8); }
    mbar_init(&sm.o_done, 1);
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
      if (elect_one()) v1_producer(sm, &tmq, &tmk, &tmv, u, bh);
    } else if (warp == 1) {
      if (elect_one()) v1_mma(sm, tmem, u);
    }
  } else {
    setmaxnreg_inc<200>();
    v3_softmax<kPolyEvery>(sm, tmem, u, p, (warp >> 2) - 1, b, h);
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
// A descriptor depends only on (pointer, B*H, rows, box): the engine launches K2 57 times per
// forward on the same few buffers, so the encoded maps are kept in a small per-thread table
// (no locks, no cross-thread state) instead of three driver calls per launch.
static ca_status make_qkv_tmap(CUtensorMap* m, const void* base, int BH, int Ntot, int box_rows) {
  struct Entry { const void* base; int bh, n, box; CUtensorMap map; };
  constexpr int kSlots = 16;
  thread_local Entry cache[kSlots];
  thread_local int used = 0, next = 0;
  for (int i = 0; i < used; ++i)
    if (cache[i].base == base && cache[i].bh == BH && cache[i].n == Ntot && cache[i].box == box_rows) {
      *m = cache[i].map;
      return CA_OK;
    }
  struct Store {
    Entry* c; int* used; int* next; const void* base; int bh, n, box; const CUtensorMap* m; bool ok = false;
    ~Store() {
      if (!ok) return;
      Entry& e = c[*next];
      e.base = base; e.bh = bh; e.n = n; e.box = box; e.map = *m;
      *next = (*next + 1) % kSlots;
      if (*used < kSlots) ++*used;
    }
  } store{cache, &used, &next, base, BH, Ntot, box_rows, m};
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
  store.ok = true;
  return CA_OK;
}

// Stream-K plan for the main units of one launch.  W units of n key steps each on S SMs: the
// first floor(W/S)*S run whole (one CTA each); the remaining R < S would leave S-R SMs idle for a
// whole unit time, so their R*n key steps are cut into S equal ranges instead (each range touches
// at most two units -> two CTAs per range, pieces of a unit merged by the last one to finish).
struct SkPlan { int n_dp, sk_units, sk_sms; };

static SkPlan plan_streamk_core(int W, int n_concept, int n_steps, int S, bool auto_rules);

// plan_sms: 0 = auto (real SM count, measured thresholds), < 0 = never split, n > 0 = plan for n SMs (tests)
static SkPlan plan_streamk(int W, int n_concept, int n_steps, bool allowed, int plan_sms) {
  SkPlan pl{W, 0, 0};
  if (!allowed || plan_sms < 0 || W == 0) return pl;
  int S = plan_sms;
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
  return plan_streamk_core(W, n_concept, n_steps, S, plan_sms == 0);
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

}  // namespace ca


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

namespace ca {
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
// Key-step ranges for the persistent v11 (see SkpIter): on when whole units would leave the last
// wave at least 8 % longer than an even split of the work, and a CTA's share is at least 8 key steps.
// Persistent v11 with a balanced tail: the flat list [concept | main] is run in whole waves of G
// CTAs in lockstep (static stride: the CTAs of a wave work on neighbouring heads, which keeps their
// K/V in L2) and only the R main units left over at the end are cut into key-step ranges (SkpIter).
// n_dp = main units run whole, R = tail units.  When the tail is tiny (a unit would be cut into more
// than 14 pieces) one whole wave is added to it.
struct SkpPlan { bool on; int G, wc; int64_t n_dp, R; };
static SkpPlan plan_v11p(int B, int H, int P, int T, int C, int mode, bool force = false) {
  SkpPlan pl{false, 0, 0, 0, 0};
  const int sms = sm_count_cached();
  if (sms <= 0 || B <= 0 || H <= 0 || P + T <= 0) return pl;
  const int n = (P + T + kBM - 1) / kBM;
  const int64_t W = (int64_t)((n + 1) / 2) * B * H;
  const int64_t n_concept = (C > 0 && mode != CA_ATTN_NEITHER) ? (int64_t)B * H : 0;
  int nc = 0;                                              // key tiles of a concept unit
  if (n_concept) {
    if (mode != CA_ATTN_SELF_ONLY) nc += (P + kBN - 1) / kBN;
    if (mode != CA_ATTN_CROSS_ONLY) nc += (C + kBN - 1) / kBN;
  }
  pl.G = sms; pl.wc = (2 * nc + 2) / 3;                    // one tile per step instead of two: ~2/3 of the cost
  const int64_t E = W + n_concept;
  int64_t e_whole = (E / sms) * sms;                       // entries of the flat list run in whole waves
  if ((E - e_whole) * 14 < sms && e_whole >= sms) e_whole -= sms;
  if (e_whole < n_concept) e_whole = n_concept;            // concept units are never cut
  pl.n_dp = e_whole - n_concept;
  pl.R = W - pl.n_dp;
  const int64_t T_cost = W * n + n_concept * pl.wc;
  const int64_t whole = ((E + sms - 1) / sms) * n;          // time of the launch without ranges, in key steps
  // structural limits: table size, tail share of at least 8 key steps per CTA, at most ~14 pieces per unit
  pl.on = sms <= kSkpMaxCtas && pl.R > 0 && pl.R * n >= (int64_t)8 * sms && pl.R * 14 >= sms &&
          pl.R * n < 0x7fffffffLL;
  // measured rule (skipped when forced by plan_sms > 0; profiles/r02_k2_variants.md): two partial pieces
  // and a merge cost a CTA about ten key steps, so the even split must save at least 24 (config 2
  // single blocks would save 8 and lose 8 %; config 5 saves 67-134 and gains 3-20 %)
  if (pl.on && !force) pl.on = whole * sms >= T_cost + (int64_t)24 * sms;
  return pl;
}
static size_t skp_workspace_bytes(const SkpPlan& pl) {
  return (size_t)2 * pl.G * kPartFloats * sizeof(float) + (size_t)2 * pl.R * sizeof(int) + 256;
}
}  // namespace ca

// Which kernel a launch uses (include/conceptattention_b200.h).  CA_K2_AUTO follows what was measured
// INSIDE the pipeline (bench.py with CA_K2_IMPL=..., profiles/r02_k2_variants.md), which differs from the
// isolated launches of scripts/bench_attention.py: in isolation the persistent v11 is 1-8 % ahead of one
// unit per CTA on every shape, inside a forward (other kernels before and after, SM clock ~1.45 GHz under
// the power cap) it is 2-8 % behind at configs 2-4 and at config 5 on one GPU (0.245 vs 0.225 ms, 2.95
// vs 2.88 ms) — its static unit-to-CTA assignment cannot absorb SMs that start late or run slower, the
// hardware scheduler of the one-unit-per-CTA launch can.  It stays ahead where its balanced tail removes
// a mostly empty last wave: config 5 sharded over 8 GPUs (3 heads: 0.371 vs 0.408-0.437 ms in the 8-GPU
// run) and over 2 (12 heads, isolated 1.26 vs 1.36-1.45), and for many short units (config 3).
// (Part of the first in-pipeline loss was the tail plan itself: it was on for config 2's single blocks,
// where it saves 8 key steps and costs 10; with the rule in plan_v11p the gap at config 2 is 3 %.)
static int resolve_impl(int impl, int B, int H, int P, int T, int C, int mode) {
  if (impl != CA_K2_AUTO) return impl;
  const int n_steps = (P + T + ca::kBM - 1) / ca::kBM;
  const ca::SkpPlan sp = ca::plan_v11p(B, H, P, T, C, mode);
  const int64_t entries = sp.n_dp + sp.R + ((C > 0 && mode != CA_ATTN_NEITHER) ? (int64_t)B * H : 0);
  // long units in few waves whose left-over wave is worth balancing (config 5 over 2-8 GPUs)
  if (n_steps > 64 && sp.on && entries < (int64_t)8 * sp.G) return CA_K2_V11P;
  // many short units: set-up per unit dominates (config 3 in the pipeline: 0.413 vs 0.456 ms)
  if (n_steps <= 16 && sp.G > 0 && entries >= (int64_t)4 * sp.G) return CA_K2_V11P;
  return CA_K2_V11;
}

extern "C" size_t ca_joint_attention_workspace_bytes_ex(int B, int H, int P, int T, int C, int mode,
                                                        int impl, int plan_sms) {
  using namespace ca;
  if (B <= 0 || H <= 0 || P + T <= 0) return 0;
  const int n_main = (P + T + kBM - 1) / kBM;
  const int which = resolve_impl(impl, B, H, P, T, C, mode);
  if (which == CA_K2_V11P && plan_sms >= 0) {                 // key-step ranges of the persistent kernel
    const SkpPlan sp = plan_v11p(B, H, P, T, C, mode, plan_sms > 0);
    return sp.on ? skp_workspace_bytes(sp) : 0;
  }
  if (which != CA_K2_V3) return 0;
  const int n_concept = (C > 0 && mode != CA_ATTN_NEITHER) ? B * H : 0;
  const SkPlan pl = plan_streamk(((n_main + 1) / 2) * B * H, n_concept, n_main, true, plan_sms);
  if (pl.sk_sms == 0) return 0;
  return (size_t)2 * pl.sk_sms * kPartFloats * sizeof(float) + (size_t)pl.sk_units * sizeof(int) + 256;
}

extern "C" size_t ca_joint_attention_workspace_bytes(int B, int H, int P, int T, int C, int mode) {
  return ca_joint_attention_workspace_bytes_ex(B, H, P, T, C, mode, CA_K2_AUTO, 0);
}

namespace ca {
// cudaFuncSetAttribute once per kernel and process, not once per launch
// (the kernel is a template VALUE: one `configured` flag per kernel, not per function type)
template <auto kKern> static ca_status launch_k2(int smem, int64_t ctas, cudaStream_t st,
                                                 const CUtensorMap& tmq, const CUtensorMap& tmk,
                                                 const CUtensorMap& tmv, const AttnParams& p) {
  static bool configured = false;
  if (!configured) {
    CA_CUDA_OK(cudaFuncSetAttribute(kKern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  kKern<<<(unsigned)ctas, 384, smem, st>>>(tmq, tmk, tmv, p);
  CA_LAUNCH_CHECK();
  return CA_OK;
}
}  // namespace ca

extern "C" ca_status ca_joint_attention_ex(const void* q, const void* k, const void* v, int B, int H,
                                           int P, int T, int C, int mode, void* out_img,
                                           void* out_txtc, void* workspace, size_t workspace_bytes,
                                           int impl, int plan_sms, void* stream) {
  using namespace ca;
  if (!q || !k || !v || B < 0 || H <= 0 || P < 0 || T < 0 || C < 0) return CA_ERR_INVALID_ARG;
  if (mode < CA_ATTN_BOTH || mode > CA_ATTN_NEITHER) return CA_ERR_INVALID_ARG;
  if (impl < CA_K2_AUTO || impl > CA_K2_V11P) return CA_ERR_INVALID_ARG;
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
  p.skp_on = 0; p.skp_wc = 0;

  CUtensorMap tmq, tmk, tmv;
  ca_status s;
  if ((s = make_qkv_tmap(&tmq, q, B * H, Ntot, kBM)) != CA_OK) return s;
  if ((s = make_qkv_tmap(&tmk, k, B * H, Ntot, kBN)) != CA_OK) return s;
  if ((s = make_qkv_tmap(&tmv, v, B * H, Ntot, kBN)) != CA_OK) return s;

  const bool concept_attn = C > 0 && mode != CA_ATTN_NEITHER;
  const int kernel = resolve_impl(impl, B, H, P, T, C, mode);
  const int W = p.n_pairs * B * H;
  p.n_concept = concept_attn ? B * H : 0;
  const size_t need = ca_joint_attention_workspace_bytes_ex(B, H, P, T, C, mode, impl, plan_sms);
  const SkPlan pl = plan_streamk(W, p.n_concept, p.n_main_tiles,
                                 kernel == CA_K2_V3 && workspace != nullptr && need > 0 && workspace_bytes >= need,
                                 plan_sms);
  p.n_dp = pl.n_dp; p.sk_units = pl.sk_units; p.sk_sms = pl.sk_sms;
  if (pl.sk_sms > 0) {
    p.ws_part = (float*)workspace;
    p.ws_count = (int*)((char*)workspace + (size_t)2 * pl.sk_sms * kPartFloats * sizeof(float));
    CA_CUDA_OK(cudaMemsetAsync(p.ws_count, 0, (size_t)pl.sk_units * sizeof(int), st));
  }
  const int64_t ctas = (int64_t)p.n_dp + 2 * p.sk_sms + p.n_concept;
  if (ctas > 0x7fffffffLL) return CA_ERR_UNSUPPORTED;
  if (ctas > 0) {
    // every kernel: polynomial exp2 on every 4th element pair (the MUFU / FMA-pipe balance point)
    if (kernel == CA_K2_V11P) {
      const SkpPlan sp = plan_v11p(B, H, P, T, C, mode, plan_sms > 0);
      if (sp.on && plan_sms >= 0 && workspace != nullptr && workspace_bytes >= skp_workspace_bytes(sp)) {
        p.skp_on = 1; p.skp_wc = sp.wc;
        p.n_dp = (int)sp.n_dp; p.sk_units = (int)sp.R;
        p.ws_part = (float*)workspace;
        p.ws_count = (int*)((char*)workspace + (size_t)2 * sp.G * kPartFloats * sizeof(float));
        CA_CUDA_OK(cudaMemsetAsync(p.ws_count, 0, (size_t)2 * sp.R * sizeof(int), st));
      }
      static bool configured = false;
      static int sms = 0;
      const int smem = (int)sizeof(SmemV11) + 1024;
      if (!configured) {
        int dev = 0;
        CA_CUDA_OK(cudaGetDevice(&dev));
        CA_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        CA_CUDA_OK(cudaFuncSetAttribute(attn_v11p_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        configured = true;
      }
      attn_v11p_kernel<4><<<(unsigned)(p.skp_on || ctas >= sms ? sms : ctas), 384, smem, st>>>(tmq, tmk, tmv, p, (int)ctas);
      CA_LAUNCH_CHECK();
      s = CA_OK;
    } else if (kernel == CA_K2_V11) s = launch_k2<attn_v11_kernel<4>>((int)sizeof(SmemV11) + 1024, ctas, st, tmq, tmk, tmv, p);
    else if (kernel == CA_K2_V7) s = launch_k2<attn_v7_kernel<4>>((int)sizeof(SmemV7) + 1024, ctas, st, tmq, tmk, tmv, p);
    else s = launch_k2<attn_v3_kernel<4>>((int)sizeof(SmemV1) + 1024, ctas, st, tmq, tmk, tmv, p);
    if (s != CA_OK) return s;
  }
  if (C > 0 && mode == CA_ATTN_NEITHER) {
    const int64_t total = (int64_t)B * H * C * (kD / 8);
    concept_v_copy_kernel<<<(unsigned)((total + 127) / 128), 128, 0, st>>>((const __nv_bfloat16*)v, p);
    CA_LAUNCH_CHECK();
  }
  return CA_OK;
}

extern "C" ca_status ca_joint_attention_ws(const void* q, const void* k, const void* v, int B, int H,
                                           int P, int T, int C, int mode, void* out_img,
                                           void* out_txtc, void* workspace, size_t workspace_bytes,
                                           void* stream) {
  return ca_joint_attention_ex(q, k, v, B, H, P, T, C, mode, out_img, out_txtc, workspace, workspace_bytes,
                               CA_K2_AUTO, 0, stream);
}

extern "C" ca_status ca_joint_attention(const void* q, const void* k, const void* v, int B, int H,
                                        int P, int T, int C, int mode, void* out_img,
                                        void* out_txtc, void* stream) {
  return ca_joint_attention_ex(q, k, v, B, H, P, T, C, mode, out_img, out_txtc, nullptr, 0, CA_K2_AUTO, 0, stream);
}

// ---- no-allocation variant (see include/conceptattention_b200.h) -------------
// The scratch of the tail balancing lives in a file-scope device array.  In host code its name is
// only a stand-in: the device address comes from cudaGetSymbolAddress, once per device, with that
// device current.  Shared by every call on the device; the callee clears the arrival counters on
// the stream at the start of each call.
static __device__ __align__(256) unsigned char g_k2_noalloc_scratch[CA_K2_NOALLOC_MAX_BYTES];

extern "C" ca_status ca_joint_attention_noalloc(const void* q, const void* k, const void* v, int B, int H,
                                                int P, int T, int C, int mode, void* out_img,
                                                void* out_txtc, int impl, int plan_sms, void* stream) {
  constexpr int kMaxDevices = 64;
  static void* scratch_of[kMaxDevices] = {};
  if (B < 0 || H <= 0 || P < 0 || T < 0 || C < 0) return CA_ERR_INVALID_ARG;
  const size_t need = ca_joint_attention_workspace_bytes_ex(B, H, P, T, C, mode, impl, plan_sms);
  if (need > (size_t)CA_K2_NOALLOC_MAX_BYTES) return CA_ERR_SCRATCH_LIMIT;
  int dev = 0;
  CA_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return CA_ERR_UNSUPPORTED;
  if (!scratch_of[dev]) {
    void* ptr = nullptr;
    CA_CUDA_OK(cudaGetSymbolAddress(&ptr, g_k2_noalloc_scratch));
    scratch_of[dev] = ptr;
  }
  return ca_joint_attention_ex(q, k, v, B, H, P, T, C, mode, out_img, out_txtc, scratch_of[dev],
                               (size_t)CA_K2_NOALLOC_MAX_BYTES, impl, plan_sms, stream);
}
