// K2 v7 / v11 — included by k2_attention.cu inside namespace ca (uses its tile constants, UnitV1,
// decode_unit, key_tile, exp2_poly2, use_poly, store_out_row).
//
// v7: v5's thread-per-row softmax with P handed to the tensor core through SHARED memory instead
// of over the S columns in TMEM, so that the score buffer of a query tile is free the moment the
// softmax warps have pulled S_t(j) into registers.
//
// Why (profiles/r02_k2_stalls.md, profiles/r02_umma_rate.md): in v3/v5 P_t overlays S_t, so
// S_t(j+1) = Q_t K(j+1)^T cannot be issued before O_t += P_t(j) V(j) has read P — the per-tile chain
// is softmax -> PV -> QK -> softmax and a key step costs T_softmax + 1024 tensor cycles (measured
// 3100 cycles, tensor pipe 66 % busy, softmax warps 36 % of their time on s_full).  A second S
// buffer per tile does not fit TMEM (S0 S1 O0 O1 = 512 columns) and halving the key chunk does not
// help: one tcgen05.mma costs at least 48 cycles, so eight N=64 MMAs take 384 cycles, not 256 (v6).
// Here the chain is cut instead: QK_t(j+2) is issued while softmax_t works on step j+1, the softmax
// never waits for a QK, and the tensor pipe only needs P_t(j) (two 64-key halves, one mbarrier
// each) to run PV.  The SS form of PV reads P (A operand, K-major SW128, the Q-tile layout) from
// smem; tcgen05.mma runs SS and TS at the same 64 cycles per N=128 K=16 step.
//
// Shared memory: Q 2 x 32 KB, P 2 x 32 KB, 3 K/V stages x 32 KB (224 KB).  Ring order = issue
// order: K0 K1 | V0 K2 | V1 K3 | ...  (K runs two key steps ahead of V).
// TMEM map as v1: S0 S1 O0 O1.

constexpr int kV7Stages = 3;

struct SmemV7 {
  uint8_t q[2][kTileBytes];
  uint8_t p[2][kTileBytes];
  uint8_t kv[kV7Stages][kTileBytes];
  uint64_t q_full[2], kv_full[kV7Stages], kv_empty[kV7Stages];
  uint64_t s_full[2], s_free[2], p_full[2][2], pv_done[2], o_done;
  uint32_t tmem_base;
};

// The first loads of a CTA (Q0, K0, Q1) are issued by thread 0 right after it has initialised the
// barriers, BEFORE the TMEM allocation and the CTA-wide sync: their ~1 us round trip then runs
// under the rest of the prologue instead of after it (kHead = true); the producer warp carries on
// from ring item 1 (kHead = false).
template <typename Smem, int kStages>
__device__ __forceinline__ void k2_load_tile(uint8_t* dst, const CUtensorMap* tm, uint64_t* bar, int row0, int bh) {
  mbar_expect_tx(bar, kTileBytes);
  tma_load_3d(dst, tm, bar, 0, row0, bh);
  tma_load_3d(dst + kHalfBytes, tm, bar, 64, row0, bh);
}

template <typename Smem, int kStages>
__device__ __forceinline__ void k2_first_loads(Smem& sm, const CUtensorMap* tmq, const CUtensorMap* tmk,
                                               const UnitV1& u, int bh) {
  int row0, valid; key_tile(u.ks, u.n0, u.j_begin, row0, valid);
  k2_load_tile<Smem, kStages>(sm.q[0], tmq, &sm.q_full[0], u.q_row0[0], bh);
  k2_load_tile<Smem, kStages>(sm.kv[0], tmk, &sm.kv_full[0], row0, bh);          // ring item 0, slot 0: free at start
  if (u.n_q == 2) k2_load_tile<Smem, kStages>(sm.q[1], tmq, &sm.q_full[1], u.q_row0[1], bh);
}

__device__ __forceinline__ void v7_producer(SmemV7& sm, const CUtensorMap* tmq, const CUtensorMap* tmk,
                                            const CUtensorMap* tmv, const UnitV1& u, int bh) {
  int item = 1;                                            // item 0 (K0) and the Q tiles: k2_first_loads
  auto load_kv = [&](const CUtensorMap* tm, int j) {
    int row0, valid; key_tile(u.ks, u.n0, u.j_begin + j, row0, valid);
    const int slot = item % kV7Stages;
    mbar_wait(&sm.kv_empty[slot], ((item / kV7Stages) & 1) ^ 1);
    k2_load_tile<SmemV7, kV7Stages>(sm.kv[slot], tm, &sm.kv_full[slot], row0, bh);
    ++item;
  };
  const int n = u.n_local;
  if (n > 1) load_kv(tmk, 1);
  for (int j = 0; j < n; ++j) {
    load_kv(tmv, j);
    if (j + 2 < n) load_kv(tmk, j + 2);
  }
}

// One thread issues every tcgen05.mma of the CTA in ring order.  (Splitting the issue over three
// threads — QK, PV_0, PV_1 — was tried: same time per launch, the issuing thread is not the limit.)
__device__ __forceinline__ void v7_mma(SmemV7& sm, uint32_t tmem, const UnitV1& u) {
  constexpr uint32_t idesc_qk = make_idesc_bf16(kBM, kBN, false, false);
  constexpr uint32_t idesc_pv = make_idesc_bf16(kBM, kD, false, true);
  const uint64_t dq = make_smem_desc(smem_u32(sm.q[0]), 16, 1024);
  const uint64_t dp = make_smem_desc(smem_u32(sm.p[0]), 16, 1024);              // P tile: K-major
  const uint64_t dk = make_smem_desc(smem_u32(sm.kv[0]), 16, 1024);             // K tile: K-major
  const uint64_t dv = make_smem_desc(smem_u32(sm.kv[0]), kHalfBytes, 1024);     // V tile: MN-major
  const bool two = u.n_q == 2;
  const int n = u.n_local;
  int item = 0;
  auto kv_wait = [&]() {
    const int slot = item % kV7Stages;
    mbar_wait(&sm.kv_full[slot], (item / kV7Stages) & 1);
    ++item;
    return slot;
  };
  auto qk = [&](int t, int kslot, int jj) {               // S_t(jj) = Q_t K^T
    if (jj == 0) mbar_wait(&sm.q_full[t], 0);
    else mbar_wait(&sm.s_free[t], (jj - 1) & 1);          // S_t(jj-1) is in the softmax registers
    tc_fence_after();
    const uint64_t qa = dq + (uint64_t)(t * (kTileBytes >> 4)), ka = dk + (uint64_t)(kslot * (kTileBytes >> 4));
#pragma unroll
    for (int kk = 0; kk < kD / 16; ++kk) {
      const uint32_t off = ((kk >> 2) * kHalfBytes + (kk & 3) * 32) >> 4;
      mma_ss(tmem + t * 128, qa + off, ka + off, idesc_qk, kk > 0);
    }
    mma_commit(&sm.s_full[t]);
  };
  auto pv = [&](int t, int vslot, int j) {               // O_t (+)= P_t V in two key halves
    const uint64_t pa = dp + (uint64_t)(t * (kTileBytes >> 4)), va = dv + (uint64_t)(vslot * (kTileBytes >> 4));
#pragma unroll
    for (int hf = 0; hf < 2; ++hf) {
      mbar_wait(&sm.p_full[t][hf], j & 1);
      tc_fence_after();
#pragma unroll
      for (int kk = 4 * hf; kk < 4 * hf + 4; ++kk) {
        const uint32_t off = ((kk >> 2) * kHalfBytes + (kk & 3) * 32) >> 4;
        mma_ss(tmem + 256 + t * 128, pa + off, va + (uint64_t)(kk * (2048 >> 4)), idesc_pv, j > 0 || kk > 0);
      }
    }
    mma_commit(&sm.pv_done[t]);
  };
#pragma unroll 1
  for (int pre = 0; pre < 2 && pre < n; ++pre) {           // S_t(0), S_t(1)
    const int ks = kv_wait();
    qk(0, ks, pre);
    if (two) qk(1, ks, pre);
    mma_commit(&sm.kv_empty[ks]);
  }
#pragma unroll 1
  for (int j = 0; j < n; ++j) {
    const bool has_k = j + 2 < n;
    const int vs = kv_wait();
    tc_fence_after();
    pv(0, vs, j);
    int ks = 0;
    if (has_k) { ks = kv_wait(); qk(0, ks, j + 2); }
    if (two) pv(1, vs, j);
    mma_commit(&sm.kv_empty[vs]);
    if (has_k) {
      if (two) qk(1, ks, j + 2);
      mma_commit(&sm.kv_empty[ks]);
    }
  }
  mma_commit(&sm.o_done);
}

template <int kPolyEvery>
__device__ __forceinline__ void v7_softmax(SmemV7& sm, uint32_t tmem, const UnitV1& u,
                                           const AttnParams& p, int t, int b, int h) {
  const int warp_q = (threadIdx.x >> 5) & 3, lane = threadIdx.x & 31;
  const int row = warp_q * 32 + lane;
  const uint32_t lane_off = (uint32_t)(warp_q * 32) << 16;
  const uint32_t t_s = tmem + t * 128 + lane_off, t_o = tmem + 256 + t * 128 + lane_off;
  const float scale = p.scale_log2;
  const int q_row0 = t ? u.q_row0[1] : u.q_row0[0], q_rows = t ? u.q_rows[1] : u.q_rows[0];
  // P_t row of this thread in the K-major SW128 tile: 16-byte chunk c of 64-key block hf lives at
  // hf * 16 KB + row * 128 + ((c ^ (row & 7)) << 4)
  const uint32_t p_row = smem_u32(sm.p[t]) + (uint32_t)row * 128u;
  const uint32_t p_xor = (uint32_t)(row & 7) << 4;
  // the only ragged key tiles are the last of each key segment
  const int n_local = u.n_local;
  const int n1 = u.n_tiles - u.n0;
  const int jr0 = u.n0 - 1 - u.j_begin, jr1 = u.n_tiles - 1 - u.j_begin;
  const int vr0 = u.n0 > 0 ? (u.ks.e0 - u.ks.s0) - (u.n0 - 1) * kBN : kBN;
  const int vr1 = n1 > 0 ? (u.ks.e1 - u.ks.s1) - (n1 - 1) * kBN : kBN;
  float m_used = -INFINITY, l_sum = 0.f;

#pragma unroll 1
  for (int j = 0; j < n_local; ++j) {
    const int valid = j == jr0 ? vr0 : (j == jr1 && n1 > 0) ? vr1 : kBN;
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
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(&sm.s_free[t]);             // S_t may be overwritten by QK_t(j+1)
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
    bool pv_waited = j == 0;
    if (j > 0 && __any_sync(0xffffffffu, grow)) {
      // rare (lazy threshold): rescale O_t once PV_t(j-1) has landed
      mbar_wait(&sm.pv_done[t], (j - 1) & 1);
      tc_fence_after();
      pv_waited = true;
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
#pragma unroll
      for (int c8 = 0; c8 < 8; ++c8) {
        uint32_t pk[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int e = 64 * hf + 8 * c8 + 2 * i;
          const uint64_t x2 = fma2(pk64(__uint_as_float(s[e]), __uint_as_float(s[e + 1])), scale2, nmb2);
          float e0, e1;
          if (use_poly<kPolyEvery>(c8 * 4 + i)) {
            exp2_poly2(x2, e0, e1);
          } else {
            float x0, x1;
            upk64(x2, x0, x1);
            e0 = ex2(x0); e1 = ex2(x1);
          }
          if (i & 1) rsb = add2(rsb, pk64(e0, e1)); else rsa = add2(rsa, pk64(e0, e1));
          pk[i] = pack2(e0, e1);
        }
        if (!pv_waited) {                                   // PV_t(j-1) has finished reading P_t
          mbar_wait(&sm.pv_done[t], (j - 1) & 1);
          pv_waited = true;
        }
        st_shared_v4(p_row + hf * kHalfBytes + (((uint32_t)c8 << 4) ^ p_xor), pk[0], pk[1], pk[2], pk[3]);
      }
      fence_proxy_async_smem();                             // generic-proxy P writes -> async proxy (MMA)
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

template <int kPolyEvery, int kCtlRegs = 88, int kSmxRegs = 208>
__global__ void __launch_bounds__(384, 1)
attn_v7_kernel(const __grid_constant__ CUtensorMap tmq, const __grid_constant__ CUtensorMap tmk,
               const __grid_constant__ CUtensorMap tmv, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  SmemV7& sm = *reinterpret_cast<SmemV7*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const UnitV1 u = decode_unit(p, blockIdx.x);
  if (u.n_local == 0) return;
  const int bh = u.bh, b = bh / p.H, h = bh - b * p.H;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmq); prefetch_tmap(&tmk); prefetch_tmap(&tmv);
    mbar_init(&sm.q_full[0], 1); mbar_init(&sm.q_full[1], 1);
    for (int i = 0; i < kV7Stages; ++i) { mbar_init(&sm.kv_full[i], 1); mbar_init(&sm.kv_empty[i], 1); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&sm.s_full[i], 1); mbar_init(&sm.s_free[i], 4); mbar_init(&sm.pv_done[i], 1);
      mbar_init(&sm.p_full[i][0], 4); mbar_init(&sm.p_full[i][1], 4);
    }
    mbar_init(&sm.o_done, 1);
    fence_barrier_init();
    k2_first_loads<SmemV7, kV7Stages>(sm, &tmq, &tmk, u, bh);
  }
  if (warp == 2) { tmem_alloc(&sm.tmem_base, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  if (warp < 4) {
    // register pool of a 384-thread CTA: 384 x 168 = 64512 = 128 x 88 + 256 x 208
    setmaxnreg_dec<kCtlRegs>();
    if (warp == 0) {
      if (elect_one()) v7_producer(sm, &tmq, &tmk, &tmv, u, bh);
    } else if (warp == 1) {
      if (elect_one()) v7_mma(sm, tmem, u);
    }
  } else {
    setmaxnreg_inc<kSmxRegs>();
    const int t = (warp >> 2) - 1;
    if (t < u.n_q) v7_softmax<kPolyEvery>(sm, tmem, u, p, t, b, h);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem, 512);
}

// =============================================================================
// v11: ONE score buffer shared by the two query tiles, P_t back in TMEM next to it.
//   TMEM  S [0,128) | P0 [128,192) | P1 [192,256) | O0 [256,384) | O1 [384,512)
//   smem  Q 2 x 32 KB + FIVE K/V stages (v7's two P tiles cost it all but three, and removing the
//         K/V reloads from a v7 launch makes it 9 % faster: the ring is too shallow)
// The softmax warps free S as soon as it is in their registers (v7), so the two tiles alternate on
// the one buffer: S0(j) -> load -> S1(j) -> load -> S0(j+1) ...; each S load (~300 cycles) sits
// under the other tile's PV on the tensor pipe, and PV_t(j) never gates a QK.  Issue order
//   QK0(0) QK1(0) | QK0(j+1) PV0(j) QK1(j+1) PV1(j) | ...      ring order K0 | K1 V0 | K2 V1 | ...
// The k-th QK of the CTA waits for phase k-1 of the single s_free barrier.
// =============================================================================
constexpr int kV11Stages = 5;
constexpr int kSkpMaxCtas = 256;                         // persistent CTAs the key-step ranges are planned for (<= SMs)

struct SmemV11 {
  uint8_t q[2][kTileBytes];
  uint8_t kv[kV11Stages][kTileBytes];
  uint64_t q_full[2], kv_full[kV11Stages], kv_empty[kV11Stages];
  uint64_t s_full[2], s_free, p_full[2][2], pv_done[2], o_done;
  uint64_t q_empty[2];                                      // persistent form only
  uint32_t tmem_base;
  int sk_flag[2];                                           // stream-K: "this piece merges" per tile
  int sk_n[2], sk_list[2][16];                              // slots of the pieces of the unit being merged
  int skp_bound[kSkpMaxCtas + 1];                           // stream-K: CTA i works on key steps [bound[i], bound[i+1])
};

// Phase counters that run on from unit to unit in the persistent form (all zero for a
// one-unit CTA): every mbarrier wait below takes its parity from them.
struct V11Carry {
  int item = 0;        // K/V ring items so far (producer and MMA thread each keep their own copy)
  int n_qk = 0;        // score tiles issued so far: the k-th waits for phase k-1 of s_free
  int steps[2] = {0, 0};   // key steps so far per tile: s_full / p_full / pv_done phases
  int units[2] = {0, 0};   // units so far that used tile t: q_full / q_empty phases
  int n_units = 0;     // units so far: o_done phases
};

// ---- v11p stream-K: key-step ranges ---------------------------------------------------------
// With skp_on the G persistent CTAs do not take whole main units: the W * n key steps of the W
// main units (n = key tiles per unit) are one stream and CTA c works on steps [B(c), B(c+1)) of it,
// after its share of the (whole) concept units c, c + G, ...  B(c) = running max over i <= c of
// floor(i * T / G) - C(i), with T = W * n + n_concept * wc the total cost and C(i) the concept cost
// of CTAs 0..i-1, so a CTA that also runs a concept unit gets that many fewer main steps.  A unit
// cut by a boundary is worked on by two (or more) CTAs: each piece leaves (O unnormalised, row max,
// row sum) per tile in its CTA's slot and the piece that finishes last merges — v3's fix-up, with
// the pieces of one range run back to back by ONE CTA.
struct SkpIter {                                          // incremental B(i), i = 0, 1, 2, ...
  int64_t Sm, conc, best, quo;                              // quo = floor(i * T / G), rem = (i * T) % G
  int64_t tq; int tr, rem;                                  // T = tq * G + tr
  int i, G, n_concept, n_whole, wc, n;
  // cost of the WHOLE entries of CTA i (flat list [concept | main], entries i, i + G, ... below n_whole)
  __device__ __forceinline__ int64_t whole_cost(int c) const {
    if (c >= n_whole) return 0;
    const int cnt = (n_whole - c + G - 1) / G;
    const int ccn = c < n_concept ? (n_concept - c + G - 1) / G : 0;
    return (int64_t)ccn * wc + (int64_t)(cnt - ccn) * n;
  }
  __device__ __forceinline__ void init(const AttnParams& p, int G_) {
    G = G_; n_concept = p.n_concept; wc = p.skp_wc; n = p.n_main_tiles;
    n_whole = p.n_concept + p.n_dp;
    Sm = (int64_t)p.sk_units * n;                           // key steps of the tail units
    // total cost: every concept unit wc, every whole main unit n, plus the tail
    const int64_t T = Sm + (int64_t)n_concept * wc + (int64_t)p.n_dp * n;
    tq = T / G; tr = (int)(T - tq * G);                     // the only 64-bit division
    conc = 0; best = 0; quo = 0; rem = 0; i = 0;
  }
  __device__ __forceinline__ int64_t bound() const {       // B(i)
    if (i >= G) return Sm;
    const int64_t v = quo - conc;
    const int64_t b = v > best ? v : best;
    return b < Sm ? b : Sm;
  }
  __device__ __forceinline__ void step() {                 // i -> i + 1
    const int64_t v = quo - conc;
    if (v > best) best = v;
    conc += whole_cost(i);
    quo += tq; rem += tr;
    if (rem >= G) { rem -= G; ++quo; }
    ++i;
  }
};

// tail unit a (0 <= a < sk_units; main unit n_dp + a) as a piece [j_begin, j_begin + n_local) of its
// key steps; slot < 0: the whole unit
__device__ __forceinline__ UnitV1 skp_piece(const AttnParams& p, int a, int j_begin, int n_local, int slot) {
  AttnParams q = p;
  q.n_dp = 0x7fffffff;                                     // decode as a plain main unit
  UnitV1 u = decode_unit(q, p.n_concept + p.n_dp + a);
  u.j_begin = j_begin; u.n_local = n_local; u.part_slot = slot; u.sk_unit = a;
  return u;
}

// Calls f(unit) for every piece of CTA `first` in the order every role uses.
template <typename F>
__device__ __forceinline__ void v11p_for_each_piece(const AttnParams& p, const int* bound, int first, int stride,
                                                    int total, F&& f) {
  if (!p.skp_on) {
#pragma unroll 1
    for (int cta = first; cta < total; cta += stride) {
      const UnitV1 u = decode_unit(p, cta);
      if (u.n_local > 0) f(u);
    }
    return;
  }
#pragma unroll 1
  for (int cta = first; cta < p.n_concept + p.n_dp; cta += stride) {   // whole entries, in lockstep with the other CTAs
    const UnitV1 u = decode_unit(p, cta);
    if (u.n_local > 0) f(u);
  }
  int64_t lo = bound[first], hi = bound[first + 1];         // table built once per CTA in the prologue
  const int n = p.n_main_tiles;
  bool is_first = true;
#pragma unroll 1
  while (lo < hi) {
    const int a = (int)(lo / n), j0 = (int)(lo - (int64_t)a * n);
    const int nl = (int)((hi - lo) < (int64_t)(n - j0) ? (hi - lo) : (int64_t)(n - j0));
    const bool whole = j0 == 0 && nl == n;
    f(skp_piece(p, a, j0, nl, whole ? -1 : 2 * first + (is_first ? 0 : 1)));
    lo += nl;
    is_first = false;
  }
}

__device__ __forceinline__ void v11_producer(SmemV11& sm, const CUtensorMap* tmq, const CUtensorMap* tmk,
                                             const CUtensorMap* tmv, const UnitV1& u, int bh, V11Carry& c,
                                             bool first_loads_done) {
  auto load_kv = [&](const CUtensorMap* tm, int j) {
    int row0, valid; key_tile(u.ks, u.n0, u.j_begin + j, row0, valid);
    const int slot = c.item % kV11Stages;
    mbar_wait(&sm.kv_empty[slot], ((c.item / kV11Stages) & 1) ^ 1);
    k2_load_tile<SmemV11, kV11Stages>(sm.kv[slot], tm, &sm.kv_full[slot], row0, bh);
    ++c.item;
  };
  const int n = u.n_local;
  if (first_loads_done) {
    ++c.item;                                              // item 0 (K0) and the Q tiles: k2_first_loads
  } else {
#pragma unroll
    for (int t = 0; t < 2; ++t) {
      if (t < u.n_q) {
        // the last QK of the unit before on this tile has read the old Q (persistent form)
        if (c.units[t] > 0) mbar_wait(&sm.q_empty[t], (c.units[t] - 1) & 1);
        k2_load_tile<SmemV11, kV11Stages>(sm.q[t], tmq, &sm.q_full[t], u.q_row0[t], bh);
      }
    }
    load_kv(tmk, 0);
  }
  for (int t = 0; t < u.n_q; ++t) ++c.units[t];
  for (int j = 0; j < n; ++j) {
    if (j + 1 < n) load_kv(tmk, j + 1);
    load_kv(tmv, j);
  }
}

__device__ __forceinline__ void v11_mma(SmemV11& sm, uint32_t tmem, const UnitV1& u, V11Carry& c) {
  constexpr uint32_t idesc_qk = make_idesc_bf16(kBM, kBN, false, false);
  constexpr uint32_t idesc_pv = make_idesc_bf16(kBM, kD, false, true);
  const uint64_t dq = make_smem_desc(smem_u32(sm.q[0]), 16, 1024);
  const uint64_t dk = make_smem_desc(smem_u32(sm.kv[0]), 16, 1024);             // K tile: K-major
  const uint64_t dv = make_smem_desc(smem_u32(sm.kv[0]), kHalfBytes, 1024);     // V tile: MN-major
  const bool two = u.n_q == 2;
  const int n = u.n_local;
  auto kv_wait = [&]() {
    const int slot = c.item % kV11Stages;
    mbar_wait(&sm.kv_full[slot], (c.item / kV11Stages) & 1);
    ++c.item;
    return slot;
  };
  auto qk = [&](int t, int kslot, int jj) {               // S = Q_t K(jj)^T
    if (jj == 0) mbar_wait(&sm.q_full[t], c.units[t] & 1);
    if (c.n_qk > 0) mbar_wait(&sm.s_free, (c.n_qk - 1) & 1);   // the score tile before this one is in registers
    tc_fence_after();
    const uint64_t qa = dq + (uint64_t)(t * (kTileBytes >> 4)), ka = dk + (uint64_t)(kslot * (kTileBytes >> 4));
#pragma unroll
    for (int kk = 0; kk < kD / 16; ++kk) {
      const uint32_t off = ((kk >> 2) * kHalfBytes + (kk & 3) * 32) >> 4;
      mma_ss(tmem, qa + off, ka + off, idesc_qk, kk > 0);
    }
    mma_commit(&sm.s_full[t]);
    ++c.n_qk;
    if (jj == n - 1) mma_commit(&sm.q_empty[t]);           // Q_t may be replaced (persistent form)
  };
  auto pv = [&](int t, int vslot, int j) {               // O_t (+)= P_t V in two key halves (P in TMEM)
    const uint64_t va = dv + (uint64_t)(vslot * (kTileBytes >> 4));
    const uint32_t par = (uint32_t)(c.steps[t] + j) & 1u;
#pragma unroll
    for (int hf = 0; hf < 2; ++hf) {
      mbar_wait(&sm.p_full[t][hf], par);
      tc_fence_after();
#pragma unroll
      for (int kk = 4 * hf; kk < 4 * hf + 4; ++kk)
        mma_ts(tmem + 256 + t * 128, tmem + 128 + t * 64 + kk * 8, va + (uint64_t)(kk * (2048 >> 4)), idesc_pv,
               j > 0 || kk > 0);
    }
    mma_commit(&sm.pv_done[t]);
  };
  int ks = kv_wait();
  qk(0, ks, 0);
  if (two) qk(1, ks, 0);
  mma_commit(&sm.kv_empty[ks]);
#pragma unroll 1
  for (int j = 0; j < n; ++j) {
    const bool has_k = j + 1 < n;
    if (has_k) { ks = kv_wait(); qk(0, ks, j + 1); }
    const int vs = kv_wait();
    tc_fence_after();
    pv(0, vs, j);
    if (has_k) {
      if (two) qk(1, ks, j + 1);
      mma_commit(&sm.kv_empty[ks]);
    }
    if (two) pv(1, vs, j);
    mma_commit(&sm.kv_empty[vs]);
  }
  mma_commit(&sm.o_done);
  for (int t = 0; t < u.n_q; ++t) { c.steps[t] += n; ++c.units[t]; }
  ++c.n_units;
}

// Stream-K piece of a main unit (u.part_slot >= 0): leave (O unnormalised, row max, row sum) of tile t
// in this piece's slot; the piece of (unit, tile) that arrives last merges all of them in range
// order — fixed, so the result does not depend on who finishes last — and stores bf16.
__device__ __forceinline__ void v11_partial_epilogue(SmemV11& sm, const AttnParams& p, const UnitV1& u, int t,
                                                     int row, uint32_t t_o, float m_used, float l_sum,
                                                     int b, int h, int q_row0, int q_rows) {
  float* mine = p.ws_part + (size_t)u.part_slot * kPartFloats + t * kPartTile;
  mine[kBM * kD + row] = m_used;
  mine[kBM * kD + kBM + row] = l_sum;
#pragma unroll
  for (int c = 0; c < kD; c += 32) {
    uint32_t r[32];
    tmem_ld32(t_o + c, r);
    tmem_wait_ld32(r);
#pragma unroll
    for (int i = 0; i < 32; i += 4)
      *reinterpret_cast<uint4*>(mine + (((c + i) >> 2) * kBM + row) * 4) =
          make_uint4(r[i], r[i + 1], r[i + 2], r[i + 3]);   // [col/4][row][4]: a warp stores 512 B runs
  }
  const int G = (int)gridDim.x, n = p.n_main_tiles;
  const int64_t u_lo = (int64_t)u.sk_unit * n, u_hi = u_lo + n;
  __threadfence();
  asm volatile("bar.sync %0, 128;" ::"r"(1 + t) : "memory");
  if ((threadIdx.x & 127) == 0) {
    // the pieces of this unit, in range order: CTA i holds one iff its range meets [u_lo, u_hi);
    // it is CTA i's first piece (slot 2i) when the range starts inside the unit, else its last (2i+1)
    int n_pieces = 0;
    const int* bnd = sm.skp_bound;
    int i0 = (int)blockIdx.x;
    while (i0 > 0 && bnd[i0] > u_lo) --i0;                // first CTA whose range reaches into the unit
    for (int i = i0; i < G && bnd[i] < u_hi; ++i) {
      if (bnd[i + 1] > bnd[i] && bnd[i + 1] > u_lo) {
        if (n_pieces < 16) sm.sk_list[t][n_pieces] = 2 * i + (bnd[i] >= u_lo ? 0 : 1);
        ++n_pieces;
      }
    }
    sm.sk_n[t] = n_pieces;
    const int ticket = atomicAdd(p.ws_count + 2 * u.sk_unit + t, 1);
    sm.sk_flag[t] = ticket == n_pieces - 1;
  }
  asm volatile("bar.sync %0, 128;" ::"r"(1 + t) : "memory");
  const bool last = sm.sk_flag[t] != 0;
  const int n_pieces = sm.sk_n[t];
  asm volatile("bar.sync %0, 128;" ::"r"(1 + t) : "memory");   // the flag may be rewritten by the next piece
  if (!last) return;                                            // (the list only after this warpgroup's next barrier)
  const int* slots = sm.sk_list[t];
  __threadfence();
  const float scale = p.scale_log2;
  float m_all = -INFINITY;
#pragma unroll 1
  for (int i = 0; i < n_pieces; ++i) {
    const float* pt = p.ws_part + (size_t)slots[i] * kPartFloats + t * kPartTile;
    m_all = fmaxf(m_all, __ldcg(pt + kBM * kD + row));
  }
#pragma unroll 1
  for (int c = 0; c < kD; c += 64) {                        // 16 x 16-byte loads in flight per piece
    float acc[64], l_all = 0.f;
#pragma unroll
    for (int e = 0; e < 64; ++e) acc[e] = 0.f;
#pragma unroll 1
    for (int i = 0; i < n_pieces; ++i) {
      const float* pt = p.ws_part + (size_t)slots[i] * kPartFloats + t * kPartTile;
      float4 o4[16];
#pragma unroll
      for (int e = 0; e < 16; ++e)
        o4[e] = __ldcg(reinterpret_cast<const float4*>(pt + (((c >> 2) + e) * kBM + row) * 4));
      const float f = ex2((__ldcg(pt + kBM * kD + row) - m_all) * scale);
      l_all += f * __ldcg(pt + kBM * kD + kBM + row);
#pragma unroll
      for (int e = 0; e < 16; ++e) {
        acc[4 * e] += f * o4[e].x; acc[4 * e + 1] += f * o4[e].y;
        acc[4 * e + 2] += f * o4[e].z; acc[4 * e + 3] += f * o4[e].w;
      }
    }
    if (row < q_rows) store_out_row(p, b, h, q_row0 + row, acc, 1.f / l_all, c, 64);
  }
}

template <int kPolyEvery>
__device__ __forceinline__ void v11_softmax(SmemV11& sm, uint32_t tmem, const UnitV1& u,
                                            const AttnParams& p, int t, int b, int h, int step0 = 0,
                                            int unit0 = 0) {
  const int warp_q = (threadIdx.x >> 5) & 3, lane = threadIdx.x & 31;
  const int row = warp_q * 32 + lane;
  const uint32_t lane_off = (uint32_t)(warp_q * 32) << 16;
  const uint32_t t_s = tmem + lane_off, t_p = tmem + 128 + t * 64 + lane_off;
  const uint32_t t_o = tmem + 256 + t * 128 + lane_off;
  const float scale = p.scale_log2;
  const int q_row0 = t ? u.q_row0[1] : u.q_row0[0], q_rows = t ? u.q_rows[1] : u.q_rows[0];
  const int n_local = u.n_local;
  const int n1 = u.n_tiles - u.n0;
  const int jr0 = u.n0 - 1 - u.j_begin, jr1 = u.n_tiles - 1 - u.j_begin;
  const int vr0 = u.n0 > 0 ? (u.ks.e0 - u.ks.s0) - (u.n0 - 1) * kBN : kBN;
  const int vr1 = n1 > 0 ? (u.ks.e1 - u.ks.s1) - (n1 - 1) * kBN : kBN;
  float m_used = -INFINITY, l_sum = 0.f;

#pragma unroll 1
  for (int j = 0; j < n_local; ++j) {
    const int valid = j == jr0 ? vr0 : (j == jr1 && n1 > 0) ? vr1 : kBN;
    mbar_wait(&sm.s_full[t], (step0 + j) & 1);
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
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(&sm.s_free);                // the score buffer goes to the other tile
    if (valid < kBN) {
#pragma unroll
      for (int i = 0; i < kBN; ++i) if (i >= valid) s[i] = 0xff800000u;   // -inf
    }
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
    const bool grow = (mx - m_used) * scale > kRescaleThreshold;
    const float m_new = grow ? mx : m_used;
    bool pv_waited = step0 + j == 0;
    if (j > 0 && __any_sync(0xffffffffu, grow)) {
      mbar_wait(&sm.pv_done[t], (step0 + j - 1) & 1);
      tc_fence_after();
      pv_waited = true;
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
      if (!pv_waited) {                                     // PV_t(j-1) has finished reading P_t
        mbar_wait(&sm.pv_done[t], (step0 + j - 1) & 1);
        tc_fence_after();
        pv_waited = true;
      }
      tmem_st32(t_p + hf * 32, pk);                         // keys [64 hf, 64 hf + 64) of P_t
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

  mbar_wait(&sm.o_done, unit0 & 1);
  tc_fence_after();
  if (u.part_slot >= 0) {
    v11_partial_epilogue(sm, p, u, t, row, t_o, m_used, l_sum, b, h, q_row0, q_rows);
    return;
  }
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

template <int kPolyEvery, int kCtlRegs = 104, int kSmxRegs = 200>
__global__ void __launch_bounds__(384, 1)
attn_v11_kernel(const __grid_constant__ CUtensorMap tmq, const __grid_constant__ CUtensorMap tmk,
                const __grid_constant__ CUtensorMap tmv, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  SmemV11& sm = *reinterpret_cast<SmemV11*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const UnitV1 u = decode_unit(p, blockIdx.x);
  if (u.n_local == 0) return;
  const int bh = u.bh, b = bh / p.H, h = bh - b * p.H;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmq); prefetch_tmap(&tmk); prefetch_tmap(&tmv);
    mbar_init(&sm.q_full[0], 1); mbar_init(&sm.q_full[1], 1);
    for (int i = 0; i < kV11Stages; ++i) { mbar_init(&sm.kv_full[i], 1); mbar_init(&sm.kv_empty[i], 1); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&sm.s_full[i], 1); mbar_init(&sm.pv_done[i], 1);
      mbar_init(&sm.p_full[i][0], 4); mbar_init(&sm.p_full[i][1], 4);
    }
    mbar_init(&sm.s_free, 4);
    mbar_init(&sm.o_done, 1);
    mbar_init(&sm.q_empty[0], 1); mbar_init(&sm.q_empty[1], 1);
    fence_barrier_init();
    k2_first_loads<SmemV11, kV11Stages>(sm, &tmq, &tmk, u, bh);
  }
  if (warp == 2) { tmem_alloc(&sm.tmem_base, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  if (warp < 4) {
    setmaxnreg_dec<kCtlRegs>();
    if (warp == 0) {
      if (elect_one()) { V11Carry c; v11_producer(sm, &tmq, &tmk, &tmv, u, bh, c, true); }
    } else if (warp == 1) {
      if (elect_one()) { V11Carry c; v11_mma(sm, tmem, u, c); }
    }
  } else {
    setmaxnreg_inc<kSmxRegs>();
    const int t = (warp >> 2) - 1;
    if (t < u.n_q) v11_softmax<kPolyEvery>(sm, tmem, u, p, t, b, h);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem, 512);
}

// =============================================================================
// v11 persistent (CA_K2_V11P): one CTA per SM walks the flat unit list with a static stride and
// calls the SAME role functions as v11 once per unit; only the mbarrier phase counters (V11Carry)
// run on.  No extra synchronisation between units: every role simply runs ahead as far as its
// barriers allow — the producer fetches the next unit's Q tiles (after q_empty: the last QK of the
// unit before has read the old ones) and K/V while the last key steps are still in flight, the MMA
// thread issues the next unit's first QK while the softmax warps drain O, and the next unit's first
// PV cannot touch an unread accumulator because it waits for a p_full that the softmax warps arrive
// on only after their epilogue (program order).  TMEM allocation, barrier init, tensor-map
// prefetch and the CTA launch itself are paid once per SM instead of once per unit.
// =============================================================================
template <int kPolyEvery>
__global__ void __launch_bounds__(384, 1)
attn_v11p_kernel(const __grid_constant__ CUtensorMap tmq, const __grid_constant__ CUtensorMap tmk,
                 const __grid_constant__ CUtensorMap tmv, const AttnParams p, int total) {
  extern __shared__ uint8_t smem_raw[];
  SmemV11& sm = *reinterpret_cast<SmemV11*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int first = (int)blockIdx.x, stride = (int)gridDim.x;
  // the very first loads go out before the CTA sync when the first piece is simply unit `first`
  // (whole units, or a concept unit ahead of the key-step range)
  const bool head_done = (!p.skp_on || first < p.n_concept + p.n_dp) && decode_unit(p, first).n_local > 0;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmq); prefetch_tmap(&tmk); prefetch_tmap(&tmv);
    for (int i = 0; i < kV11Stages; ++i) { mbar_init(&sm.kv_full[i], 1); mbar_init(&sm.kv_empty[i], 1); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&sm.q_full[i], 1); mbar_init(&sm.q_empty[i], 1);
      mbar_init(&sm.s_full[i], 1); mbar_init(&sm.pv_done[i], 1);
      mbar_init(&sm.p_full[i][0], 4); mbar_init(&sm.p_full[i][1], 4);
    }
    mbar_init(&sm.s_free, 4);
    mbar_init(&sm.o_done, 1);
    fence_barrier_init();
    if (head_done) {
      const UnitV1 u0 = decode_unit(p, first);
      k2_first_loads<SmemV11, kV11Stages>(sm, &tmq, &tmk, u0, u0.bh);
    }
  }
  if (warp == 2) { tmem_alloc(&sm.tmem_base, 512); tmem_relinquish(); }
  if (p.skp_on && threadIdx.x == 96) {                      // idle warp 3: the range table, once per CTA
    SkpIter it; it.init(p, stride);
    for (int i = 0; i <= stride; ++i) { sm.skp_bound[i] = (int)it.bound(); it.step(); }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  if (warp < 4) {
    setmaxnreg_dec<104>();
    if (warp == 0) {
      if (elect_one()) {
        V11Carry c;
        bool head = head_done;
        v11p_for_each_piece(p, sm.skp_bound, first, stride, total, [&](const UnitV1& u) {
          v11_producer(sm, &tmq, &tmk, &tmv, u, u.bh, c, head);
          head = false;
        });
      }
    } else if (warp == 1) {
      if (elect_one()) {
        V11Carry c;
        v11p_for_each_piece(p, sm.skp_bound, first, stride, total, [&](const UnitV1& u) { v11_mma(sm, tmem, u, c); });
      }
    }
  } else {
    setmaxnreg_inc<200>();
    const int t = (warp >> 2) - 1;
    int step0 = 0, unit0 = 0;
    v11p_for_each_piece(p, sm.skp_bound, first, stride, total, [&](const UnitV1& u) {
      if (t < u.n_q) {
        const int b = u.bh / p.H, h = u.bh - b * p.H;
        v11_softmax<kPolyEvery>(sm, tmem, u, p, t, b, h, step0, unit0);
        step0 += u.n_local;
      }
      ++unit0;
    });
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem, 512);
}
