// tcgen05 TF32 GEMM for throughput-bound (large-batch) learner steps: the "stream" kernel.
//
// Same problem interface and the same smem operand layouts as the slab kernel (gemm_tf32.cuh), but
// built for many tiles per SM instead of one short tile per CTA:
//   * persistent CTAs (one per SM) walk work items (z, tile_m, tile_n, split) blockIdx.x, +gridDim.x, ...
//   * operands flow through an S-stage ring of one 32-float k-block each: warps 0-7 keep S/2 stages of
//     16-byte cp.async in flight (full/empty mbarriers with phase parity), so K is unbounded and the
//     loads of the next tile overlap the MMAs of this one; uint8 operands (conv1) are prefetched RD
//     k-blocks ahead through a rotating register file and widened on their way into the ring;
//   * problems with B_RESIDENT_KB (conv weights) keep their whole B operand in shared memory instead;
//   * warp 12's elected lane issues the MMAs and releases stages with tcgen05.commit;
//   * the accumulator is double-buffered in TMEM (2 x BN columns): warps 8-11 drain tile i
//     (tcgen05.ld -> p.store4 straight from registers, thread = accumulator row) while the tensor
//     core works on tile i+1.
// Split-K exists only for DEFER_REDUCE problems (weight gradients: few tiles, K = batch x positions);
// each split is just another work item that leaves its partial through p.store4.
// grid = min(items, SMs), block = 416 threads, no clusters.
#pragma once
#include "gemm_tf32.cuh"

namespace dqnb {

constexpr int kT32SThreads = 416;   // 8 producer warps, 4 epilogue warps, 1 MMA warp
constexpr int kT32SEpiWarp = 8;     // warps 8..11: warp % 4 = the TMEM lane quarter it may read
constexpr int kT32SMmaWarp = 12;

// Optional member of a problem type:
//   static constexpr int B_RESIDENT_KB = n;   the whole B operand of a (z, tile_n) is at most n k-blocks
// (conv weights).  The stream kernel then keeps B in shared memory across tiles - it is staged once per
// (z, tile_n) a CTA meets, i.e. a few times per launch - and the ring carries A only: a third less
// staging work and shared-memory fill per k-block for the convolutions.
template <class P, class = void> struct t32_b_resident : std::integral_constant<int, 0> {};
// Filters above kT32SResidentMax would leave the ring two or three k-blocks of A (conv2 / conv3 forward,
// conv3 dgrad: 128-144 KB): those stream B with A instead, which buys two k-blocks per stage and a deeper ring.
constexpr int kT32SResidentMax = 96 * 1024;
template <class P>
struct t32_b_resident<P, std::enable_if_t<(P::B_RESIDENT_KB > 0 && P::B_RESIDENT_KB * t32_tile_bytes(P::BN, P::B_MN_MAJOR) <= kT32SResidentMax)>>
    : std::integral_constant<int, P::B_RESIDENT_KB> {};

//   static constexpr bool B_REG = true;  float4 b_get(const BRow&, const BCol&) const;
// (only with A_REG, and only when a producer thread owns a single B chunk per k-block): B also travels
// through the register prefetch.  A kernel whose operands all avoid cp.async never executes
// cp.async.wait_group - ptxas tracks plain global loads on the same scoreboard, so a wait_group in the
// loop also waits for the register prefetch and throws its depth away.
template <class P, class = void> struct t32_b_reg : std::false_type {};
template <class P> struct t32_b_reg<P, std::enable_if_t<P::B_REG>> : std::true_type {};

// k-blocks per ring stage.  A stage costs every producer warp a fixed serial latency (~500 clocks of
// barrier wait, address arithmetic, MEMBAR and arrive) whatever it carries, and the MMA thread one
// wait + commit: two k-blocks per stage halve that overhead per unit of K.  Work items are cut at
// multiples of that many k-blocks (loaders zero-fill k >= K).
// Only where the ring keeps at least three such stages: problems whose resident B leaves little shared
// memory (conv2 / conv3 forward) stay at one k-block per stage.
template <class P>
__host__ __device__ constexpr int t32s_resident_bytes() { return t32_b_resident<P>::value * t32_tile_bytes(P::BN, P::B_MN_MAJOR); }
// A_ROW16 problems (uint8 rows of 32 source bytes per k-block): the raw bytes are prefetched with
// cp.async into a byte ring of kT32SByteDepth k-blocks x 4 KB behind the operand ring; thread t reads
// back the 16 bytes it copied itself, widens them and stores the tf32 chunks.  cp.async groups retire
// in order (wait_group counts), which plain LDG prefetches do not: eight loads in flight share the
// warp's six scoreboards, so waiting for the oldest also waited for the newest.
// A_COL16 problems (conv1 weight gradient: MN-major uint8 A whose 4 consecutive mn-chunks of one k are 16
// contiguous source bytes, plus one fp32 B chunk per thread) send both operands through the same ring:
//   static constexpr bool A_COL16 = true;
//   const uint8_t* a_src16(int m0, int k, int c16) const;   16 bytes of mn-chunks 4*c16..4*c16+3 at k, or null
//   const float* b_src(const BRow&, const BCol&) const;     the thread's B chunk, or null
constexpr int kT32SByteDepth = 8;
template <class P, class = void> struct t32_a_col16 : std::false_type {};
template <class P> struct t32_a_col16<P, std::enable_if_t<P::A_COL16>> : std::true_type {};
template <class P>
__host__ __device__ constexpr int t32s_byte_slot() {
  return t32_a_row16<P>::value ? 4096 : (t32_a_col16<P>::value || t32_a_row16h<P>::value) ? 8192 : 0;
}
template <class P>
__host__ __device__ constexpr int t32s_byte_ring() { return kT32SByteDepth * t32s_byte_slot<P>(); }
template <class P>
__host__ __device__ constexpr int t32s_kb_bytes() {
  return t32_tile_bytes(kT32BM, P::A_MN_MAJOR) + (t32_b_resident<P>::value ? 0 : t32_tile_bytes(P::BN, P::B_MN_MAJOR));
}
template <class P>
__host__ __device__ constexpr int t32s_kps() {
  return (208 * 1024 - t32s_resident_bytes<P>() - t32s_byte_ring<P>()) / (2 * t32s_kb_bytes<P>()) >= 3 ? 2 : 1;
}
template <class P>
__host__ __device__ constexpr int t32s_stage_bytes() { return t32s_kps<P>() * t32s_kb_bytes<P>(); }
template <class P>
__host__ __device__ constexpr int t32s_stages() {
  return (208 * 1024 - t32s_resident_bytes<P>() - t32s_byte_ring<P>()) / t32s_stage_bytes<P>() > 8
             ? 8
             : (208 * 1024 - t32s_resident_bytes<P>() - t32s_byte_ring<P>()) / t32s_stage_bytes<P>();
}
template <class P>
__host__ __device__ constexpr int t32s_smem_bytes() {
  return t32s_resident_bytes<P>() + t32s_stages<P>() * t32s_stage_bytes<P>() + t32s_byte_ring<P>() + 1024;
}

// Timing knobs and the event trace exist only in builds that define DQNB_STREAM_KNOBS (the micro
// harnesses): every knob is a load + branch on the serial path of a ring stage.
#ifdef DQNB_STREAM_KNOBS
#define T32S_KNOB(w, bit) ((w).dbg & (bit))
#else
#define T32S_KNOB(w, bit) 0
#endif
// Division by a launch constant as multiply-high + shifts (Granlund-Montgomery): decoding a work item
// sits on the serial path of every tile boundary in all three roles, and a runtime 32-bit division is
// a ~100-instruction dependent chain.
struct T32FastDiv {
  uint32_t d, m, s1, s2;
  __device__ __forceinline__ uint32_t div(uint32_t n) const {
    const uint32_t t = __umulhi(m, n);
    return (t + ((n - t) >> s1)) >> s2;
  }
};
inline T32FastDiv t32s_fastdiv(uint32_t d) {
  uint32_t l = 0;
  while ((1ull << l) < d) ++l;
  T32FastDiv f;
  f.d = d;
  f.m = (uint32_t)(((1ull << 32) * ((1ull << l) - d)) / d + 1);
  f.s1 = l < 1 ? l : 1;
  f.s2 = l > 0 ? l - 1 : 0;
  return f;
}
struct T32Work {
  T32FastDiv fd_splits, fd_tn, fd_tm;
  int tiles_m, tiles_n, splits, per, kblocks, items;   // per = k-blocks per split; items = Z*tiles*splits
  int dbg;   // timing experiments only (DQNB_STREAM_DBG): 1 = no epilogue stores, 2 = no A st.shared, 4 = no A loads
};
struct T32Item {
  int z, split, m0, n0, kb_beg, nkb;
};
template <int BN>
__device__ __forceinline__ T32Item t32s_item(const T32Work& w, int item) {
  uint32_t r = w.fd_splits.div((uint32_t)item);
  const int split = item - (int)r * w.splits;
  uint32_t r2 = w.fd_tn.div(r);
  const int tn = (int)(r - r2 * (uint32_t)w.tiles_n);
  const int z = (int)w.fd_tm.div(r2);
  const int tm = (int)r2 - z * w.tiles_m;
  const int kb_beg = split * w.per, kb_end = min(w.kblocks, kb_beg + w.per);   // both multiples of the problem's k-blocks per stage
  return {z, split, tm * kT32BM, tn * BN, kb_beg, kb_end - kb_beg};
}

// Optional event trace (scripts/micro/stream_trace.cu): CTA 0 writes clock64() of event `idx` of series
// `series` to sk.dbg[series * 256 + idx].  sk.dbg is null in the product.
__device__ __forceinline__ void t32s_trace(long long* trace, int series, uint32_t idx) {
#ifndef DQNB_STREAM_KNOBS
  return;
#endif
  if (trace != nullptr && blockIdx.x == 0 && idx < 256) trace[series * 256 + idx] = clock64();
}

template <class P>
__global__ void __launch_bounds__(kT32SThreads, 1) gemm_tf32_stream(P p, SplitK sk, T32Work w) {
  constexpr int BN = P::BN, COLS = t32_tmem_cols(BN), S = t32s_stages<P>();
  // D = cp.async stages a producer keeps in flight before it publishes the oldest.  Stage q is handed to
  // the tensor core while stage q+D is being issued, which needs slot (q+D-S) back: the ring therefore
  // holds S-D published stages of slack between producers and MMAs (D = S-1 would lock-step them).
  // The register path prefetches A in registers; only the L2-resident B operand rides on cp.async.
  constexpr int RKB = t32_b_resident<P>::value;
  constexpr bool B_REG = t32_b_reg<P>::value;
  constexpr bool NO_CP = t32_a_reg<P>::value && (RKB > 0 || B_REG);   // no cp.async in the steady state
  static_assert(!B_REG || t32_a_reg<P>::value, "B_REG needs the register path (A_REG)");
  constexpr int D = t32_a_reg<P>::value ? (NO_CP ? 0 : 2) : S / 2;
  constexpr int A_TILE = t32_tile_bytes(kT32BM, P::A_MN_MAJOR), B_TILE = t32_tile_bytes(BN, P::B_MN_MAJOR);
  constexpr int STAGE = t32s_stage_bytes<P>(), RING = t32s_resident_bytes<P>();   // ring starts behind resident B
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t bars[2 * S + 5];   // full[S], empty[S], acc_full[2], acc_empty[2], b_full
  __shared__ uint32_t tmem_slot;

  const int t = threadIdx.x, warp = t >> 5;
  const unsigned long long tl_t0 = tl_begin(sk.tl);
  const uint32_t smem = (smem_u32(smem_raw) + 1023u) & ~1023u;
  if (t == 0) {
    for (int s = 0; s < S; ++s) { mbar_init(&bars[s], kT32Producers / 32); mbar_init(&bars[S + s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&bars[2 * S + a], 1); mbar_init(&bars[2 * S + 2 + a], 4); }
    mbar_init(&bars[2 * S + 4], kT32Producers / 32);
    fence_mbar_init();
  }
  if (warp == kT32SMmaWarp) t32_alloc<2 * COLS>(&tmem_slot);
  t32_fence_before();
  __syncthreads();
  t32_fence_after();
  const uint32_t tmem = tmem_slot;
  // shared-space addresses of the barriers, computed once
  const uint32_t full_bar = smem_u32(bars), empty_bar = full_bar + 8 * S, acc_full = full_bar + 16 * S,
                 acc_empty = acc_full + 16, b_full = acc_full + 32;

  if (warp < 8) {
    // ------------------------------------------------------------------ producers
    using MA = T32Map<kT32BM, P::A_MN_MAJOR>;
    using MB = T32Map<BN, P::B_MN_MAJOR>;
    uint32_t q = 0, pub = 0;   // stages issued / published so far by this CTA (same sequence in every role)
    int b_key = -1;            // (z, tile_n) whose B operand is resident
    typename P::BRow brow[MB::NROW];
    auto put_b = [&](const T32Item& it, int kb, uint32_t tile) {   // k-block kb of B -> the smem tile at `tile`
      const int k0 = kb * kT32BK;
#pragma unroll
      for (int uc = 0; uc < MB::NCOL; ++uc) {
        typename P::BCol col = p.b_col(it.z, k0 + MB::col_index(t, uc));
#pragma unroll
        for (int ur = 0; ur < MB::NROW; ++ur) p.b_put(brow[ur], col, tile + MB::offset(t, ur, uc));
      }
    };
    constexpr int KPS = t32s_kps<P>();
    auto stage_b = [&](const T32Item& it, int i, uint32_t base, int sub) {   // k-block i of the item = sub-block `sub` of its stage
      if constexpr (RKB == 0 && !B_REG) put_b(it, it.kb_beg + i, base + KPS * A_TILE + sub * B_TILE);
    };
    auto acquire = [&]() -> uint32_t {   // the ring slot of stage q, once the tensor core has released it
      const uint32_t s = q % S;
      if (q >= (uint32_t)S) mbar_wait_a(empty_bar + 8 * s, ((q / S) - 1) & 1);
      if (t == 0) t32s_trace(sk.dbg, 0, q);   // series 0: producer holds slot of stage q
      return smem + RING + s * STAGE;
    };
    auto publish = [&]() {   // stage q issued: hand the oldest unpublished stage (landed by now) to the tensor core
      if constexpr (NO_CP) {   // everything was written with st.shared: publish at once
        if (!T32S_KNOB(w, 8)) fence_proxy_async_smem();
        mbar_arrive_warp_a(full_bar + 8 * (q % S));
        pub = ++q;
        return;
      }
      cp_commit();
      ++q;
      if (q - pub > (uint32_t)D) {
        cp_wait<D>();
        fence_proxy_async_smem();
        mbar_arrive_warp_a(full_bar + 8 * (pub % S));
        ++pub;
      }
    };
    auto flush = [&]() {   // publish everything issued
      cp_wait<0>();
      fence_proxy_async_smem();
      for (; pub < q; ++pub) mbar_arrive_warp_a(full_bar + 8 * (pub % S));
    };
    // (brow must describe `it`.)  Resident mode: make B of this item's (z, tile_n) the resident one.
    auto open_b = [&](const T32Item& it) {
      if constexpr (RKB > 0) {
        const int key = it.z * w.tiles_n + it.n0 / BN;
        if (key == b_key) return;
        b_key = key;
        flush();
        if (q > 0) mbar_wait_a(empty_bar + 8 * ((q - 1) % S), ((q - 1) / S) & 1);   // every MMA on the old B has retired
        for (int kb = 0; kb < w.kblocks; ++kb) put_b(it, kb, smem + kb * B_TILE);
        cp_commit();
        cp_wait<0>();
        fence_proxy_async_smem();
        mbar_arrive_warp_a(b_full);
      }
    };
    if constexpr (t32_a_reg<P>::value) {
      // uint8 A through registers: the load cursor runs RD k-blocks ahead of the store cursor
      // ROW16H (fp16 operands): a k-block is 64 halves = two 32-byte patch rows; both 16-byte pieces of this
      // thread's half ride one ring slot (piece s at +4096 s)
      constexpr bool ROW16 = t32_a_row16<P>::value, COL16 = t32_a_col16<P>::value, ROW16H = t32_a_row16h<P>::value;
      constexpr bool BYTES = ROW16 || COL16 || ROW16H;
      constexpr int SLOT = t32s_byte_slot<P>();
      static_assert(!COL16 || (P::A_MN_MAJOR && B_REG), "A_COL16: MN-major A and a one-chunk-per-thread B");
      constexpr int NW = ROW16 ? 4 : MA::NROW * MA::NCOL;
      constexpr int RD = BYTES ? kT32SByteDepth : 6;
      const int row = t >> 1, half = t & 1;   // ROW16: this thread's half row of the tile
      const uint32_t stg = smem + RING + S * STAGE + t * 16;   // ROW16: this thread's 16 bytes of byte-ring slot 0
      uint32_t wreg[BYTES ? 1 : RD][NW];
      float4 breg[B_REG && !BYTES ? RD : 1];
      static_assert(!B_REG || MB::NROW * MB::NCOL == 1, "B_REG: one B chunk per producer thread and k-block");
      int l_item = blockIdx.x, l_i = 0;
      T32Item l_it = {};
      const uint8_t* l_rp = nullptr;
      typename P::ARow l_arow[MA::NROW];
      typename P::BRow l_brow;
      auto l_open = [&]() {
        if (l_item >= w.items) return;
        l_it = t32s_item<BN>(w, l_item);
        if constexpr (B_REG) l_brow = p.b_row(l_it.z, l_it.n0 + MB::row_index(t, 0));
        if constexpr (ROW16 || ROW16H) l_rp = p.a_row16(l_it.z, l_it.m0 + row);
        else if constexpr (COL16) {
        } else {
#pragma unroll
          for (int u = 0; u < MA::NROW; ++u) l_arow[u] = p.a_row(l_it.z, l_it.m0 + MA::row_index(t, u));
        }
      };
      auto l_next = [&](uint32_t (&dst)[NW], float4& bdst, int slot) {
        if constexpr (ROW16) {   // one cp.async group per k-block, also past the end (the wait counts stay static)
          const bool on = l_item < w.items && !T32S_KNOB(w, 4);
          const bool ok = on && l_rp != nullptr;
          cp16(stg + slot * 4096, ok ? l_rp + p.a_kb_off(l_it.kb_beg + l_i) + half * 16 : reinterpret_cast<const uint8_t*>(p.x[0]), ok);
          cp_commit();
          if (l_item < w.items && ++l_i == l_it.nkb) { l_i = 0; l_item += gridDim.x; l_open(); }
          return;
        }
        if constexpr (ROW16H) {
          const bool on = l_item < w.items && !T32S_KNOB(w, 4);
          const bool ok = on && l_rp != nullptr;
#pragma unroll
          for (int s = 0; s < 2; ++s)
            cp16(stg + slot * SLOT + s * 4096,
                 ok ? l_rp + p.a_kb_off(2 * (l_it.kb_beg + l_i) + s) + half * 16 : reinterpret_cast<const uint8_t*>(p.x[0]), ok);
          cp_commit();
          if (l_item < w.items && ++l_i == l_it.nkb) { l_i = 0; l_item += gridDim.x; l_open(); }
          return;
        }
        if constexpr (COL16) {   // A bytes and the B chunk of one k-block = one cp.async group
          const bool on = l_item < w.items && !T32S_KNOB(w, 4);
          const int k0 = (l_it.kb_beg + l_i) * kT32BK;
          const uint8_t* a = on ? p.a_src16(l_it.m0, k0 + (t >> 3), t & 7) : nullptr;
          cp16(stg + slot * SLOT, a != nullptr ? a : reinterpret_cast<const uint8_t*>(p.x), a != nullptr);
          const float* b = on ? p.b_src(l_brow, p.b_col(l_it.z, k0 + MB::col_index(t, 0))) : nullptr;
          cp16(stg + slot * SLOT + 4096, b != nullptr ? reinterpret_cast<const void*>(b) : reinterpret_cast<const void*>(p.x), b != nullptr);
          cp_commit();
          if (l_item < w.items && ++l_i == l_it.nkb) { l_i = 0; l_item += gridDim.x; l_open(); }
          return;
        }
#pragma unroll
        for (int j = 0; j < NW; ++j) dst[j] = 0u;
        if (l_item >= w.items) return;
        if (T32S_KNOB(w, 4)) { if (++l_i == l_it.nkb) { l_i = 0; l_item += gridDim.x; l_open(); } return; }
        if constexpr (B_REG) bdst = p.b_get(l_brow, p.b_col(l_it.z, (l_it.kb_beg + l_i) * kT32BK + MB::col_index(t, 0)));
        {
#pragma unroll
          for (int uc = 0; uc < MA::NCOL; ++uc) {
            typename P::ACol col = p.a_col(l_it.z, (l_it.kb_beg + l_i) * kT32BK + MA::col_index(t, uc));
#pragma unroll
            for (int ur = 0; ur < MA::NROW; ++ur) dst[uc * MA::NROW + ur] = p.a_get(l_arow[ur], col);
          }
        }
        if (++l_i == l_it.nkb) { l_i = 0; l_item += gridDim.x; l_open(); }
      };
      l_open();
#pragma unroll
      for (int u = 0; u < RD; ++u) l_next(wreg[BYTES ? 0 : u], breg[B_REG && !BYTES ? u : 0], u);
      int s_item = blockIdx.x, s_i = 0;
      T32Item s_it = {};
      auto s_open = [&]() {
        if (s_item >= w.items) return;
        s_it = t32s_item<BN>(w, s_item);
#pragma unroll
        for (int u = 0; u < MB::NROW; ++u) brow[u] = p.b_row(s_it.z, s_it.n0 + MB::row_index(t, u));
        open_b(s_it);
      };
      s_open();
      static_assert(RD % KPS == 0, "prefetch depth must be a whole number of stages");
      static_assert(!BYTES || RKB > 0 || B_REG, "byte-ring paths publish without cp.async waits: B must be resident or ride the ring");
      while (s_item < w.items) {
#pragma unroll
        for (int u0 = 0; u0 < RD; u0 += KPS) {
          if (s_item < w.items) {
            const uint32_t base = acquire();
            if constexpr (BYTES) {
              // the whole stage at once: one wait, all read-backs, then the widening stores, then the
              // next prefetches - a shorter dependent chain per stage than k-block by k-block
              cp_wait<RD - KPS>();
              uint32_t wd[KPS][4], wh[ROW16H ? KPS : 1][4];
              float4 bv[COL16 ? KPS : 1];
#pragma unroll
              for (int sub = 0; sub < KPS; ++sub) {
                const uint32_t src = stg + (u0 + sub) * SLOT;
                asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(wd[sub][0]), "=r"(wd[sub][1]), "=r"(wd[sub][2]), "=r"(wd[sub][3]) : "r"(src) : "memory");
                if constexpr (COL16) bv[sub] = ld_shared_f4(src + 4096);
                if constexpr (ROW16H)
                  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(wh[sub][0]), "=r"(wh[sub][1]), "=r"(wh[sub][2]), "=r"(wh[sub][3]) : "r"(src + 4096) : "memory");
              }
#pragma unroll
              for (int sub = 0; sub < KPS; ++sub) {
                const uint32_t abase = base + sub * A_TILE;
                if (T32S_KNOB(w, 2)) {
                } else if constexpr (ROW16) {
                  const uint32_t row_base = abase + (row >> 3) * 1024 + (row & 7) * 128;
#pragma unroll
                  for (int j = 0; j < 4; ++j) p.a_store(row_base + ((uint32_t)((half * 4 + j) ^ (row & 7)) << 4), wd[sub][j]);
                } else if constexpr (ROW16H) {
                  const uint32_t row_base = abase + (row >> 3) * 1024 + (row & 7) * 128, sw = (uint32_t)(row & 7);
                  const uint32_t c0 = half * 2, c1 = 4 + half * 2;   // chunks of patch row s = 0 / 1
                  t32_widen8h(row_base + ((c0 ^ sw) << 4), wd[sub][0], wd[sub][1]);
                  t32_widen8h(row_base + (((c0 + 1) ^ sw) << 4), wd[sub][2], wd[sub][3]);
                  t32_widen8h(row_base + ((c1 ^ sw) << 4), wh[sub][0], wh[sub][1]);
                  t32_widen8h(row_base + (((c1 + 1) ^ sw) << 4), wh[sub][2], wh[sub][3]);
                } else {
                  const int kk = t >> 3, c16 = t & 7;   // SWIZZLE_128B_BASE32B atom: see T32Map::offset
                  const uint32_t kbase = abase + (kk >> 2) * 512 + (kk & 3) * 128;
#pragma unroll
                  for (int j = 0; j < 4; ++j) {
                    const int cc = 4 * c16 + j;
                    p.a_store(kbase + (cc >> 3) * 4096 + ((((cc >> 1) & 3) ^ (kk & 3)) << 5) + ((cc & 1) << 4), wd[sub][j]);
                  }
                  st16(base + KPS * A_TILE + sub * B_TILE + MB::offset(t, 0, 0), bv[sub].x, bv[sub].y, bv[sub].z, bv[sub].w);
                }
              }
#pragma unroll
              for (int sub = 0; sub < KPS; ++sub) l_next(wreg[0], breg[0], u0 + sub);
            } else {
#pragma unroll
              for (int sub = 0; sub < KPS; ++sub) {
                const int u = u0 + sub;
                const uint32_t abase = base + sub * A_TILE;
                if (!T32S_KNOB(w, 2)) {
#pragma unroll
                  for (int uc = 0; uc < MA::NCOL; ++uc)
#pragma unroll
                    for (int ur = 0; ur < MA::NROW; ++ur) p.a_store(abase + MA::offset(t, ur, uc), wreg[u][uc * MA::NROW + ur]);
                }
                if constexpr (B_REG) {
                  const float4 bvv = breg[u];
                  st16(base + KPS * A_TILE + sub * B_TILE + MB::offset(t, 0, 0), bvv.x, bvv.y, bvv.z, bvv.w);
                }
                l_next(wreg[u], breg[B_REG ? u : 0], u);
                stage_b(s_it, s_i + sub, base, sub);
              }
            }
            publish();
            s_i += KPS;
            if (s_i == s_it.nkb) { s_i = 0; s_item += gridDim.x; s_open(); }
          }
        }
      }
    } else {
      typename P::ARow arow[MA::NROW];
      for (int item = blockIdx.x; item < w.items; item += gridDim.x) {
        const T32Item it = t32s_item<BN>(w, item);
#pragma unroll
        for (int u = 0; u < MA::NROW; ++u) arow[u] = p.a_row(it.z, it.m0 + MA::row_index(t, u));
#pragma unroll
        for (int u = 0; u < MB::NROW; ++u) brow[u] = p.b_row(it.z, it.n0 + MB::row_index(t, u));
        open_b(it);
#pragma unroll 1
        for (int i = 0; i < it.nkb; i += KPS) {
          const uint32_t base = acquire();
#pragma unroll
          for (int sub = 0; sub < KPS; ++sub) {
            const int k0 = (it.kb_beg + i + sub) * kT32BK;
#pragma unroll
            for (int uc = 0; uc < MA::NCOL; ++uc) {
              typename P::ACol col = p.a_col(it.z, k0 + MA::col_index(t, uc));
#pragma unroll
              for (int ur = 0; ur < MA::NROW; ++ur) p.a_put(arow[ur], col, base + sub * A_TILE + MA::offset(t, ur, uc));
            }
            stage_b(it, i + sub, base, sub);
          }
          publish();
        }
      }
    }
    flush();   // the last D stages
  } else if (warp == kT32SMmaWarp) {
    // ------------------------------------------------------------------ MMA issuer
    // The whole warp walks the barriers and one ELECTED lane issues everything the tensor core does (elect.sync keeps
    // the descriptors in uniform registers; a threadIdx test put a R2UR waterfall loop - ~10 instructions and ~96
    // clocks - around every tcgen05.mma).  This loop is the serial spine of the CTA and
    // is kept to a handful of instructions per k-block: descriptors are a constant high word plus a low
    // word that is advanced by adds (smem address >> 4 lives in bits 0-13, and the CTA's window is
    // < 256 KB, so no carry leaves the field); ring slot and phase are counters, not divisions.
    const bool leader = elect_one();
    const uint32_t idesc = t32_f16<P>::value ? t32_idesc_f16(BN) : t32_idesc(BN, P::A_MN_MAJOR, P::B_MN_MAJOR);
    const uint64_t a_d0 = P::A_MN_MAJOR ? t32_desc(smem + RING, 4096, 512, 1) : t32_desc(smem + RING, 16, 1024);
    const uint64_t b_d0 = P::B_MN_MAJOR ? t32_desc(RKB > 0 ? smem : smem + RING + t32s_kps<P>() * A_TILE, 4096, 512, 1)
                                        : t32_desc(RKB > 0 ? smem : smem + RING + t32s_kps<P>() * A_TILE, 16, 1024);
    const uint32_t a_hi = (uint32_t)(a_d0 >> 32), b_hi = (uint32_t)(b_d0 >> 32);
    constexpr uint32_t A_KS = (P::A_MN_MAJOR ? 1024 : 32) >> 4, B_KS = (P::B_MN_MAJOR ? 1024 : 32) >> 4;
    uint32_t s = 0, ph = 0, n = 0, nb = 0;   // ring slot + its phase / tile / resident-B sequence numbers
    int key_end = 0;                          // resident B: first item of the next (z) group
    const int nkb_all = w.kblocks, fast = w.splits == 1;
    for (int item = blockIdx.x; item < w.items; item += gridDim.x, ++n) {
      int nkb = nkb_all, kb_beg = 0;
      if (!fast) {
        const T32Item it = t32s_item<BN>(w, item);
        nkb = it.nkb; kb_beg = it.kb_beg;
      }
      if constexpr (RKB > 0) {   // (tiles_n == 1, splits == 1: the key is z = item / tiles_m)
        if (item >= key_end) {
          key_end = ((int)w.fd_tm.div((uint32_t)item) + 1) * w.tiles_m;
          mbar_wait_a(b_full, nb & 1);
          ++nb;
        }
      }
      const uint32_t a = n & 1;
      if (n >= 2) mbar_wait_a(acc_empty + 8 * a, ((n >> 1) - 1) & 1);   // the epilogue has drained this accumulator
      t32_fence_after();
      t32s_trace(sk.dbg, 1, n);   // series 1: accumulator of tile n available
      const uint32_t d_tmem = tmem + a * COLS;
      uint32_t b_lo_res = (uint32_t)b_d0 + (uint32_t)kb_beg * (B_TILE >> 4);
#pragma unroll 1
      for (int i = 0; i < nkb; i += t32s_kps<P>()) {
        mbar_wait_a(full_bar + 8 * s, ph);
        t32_fence_after();
        t32s_trace(sk.dbg, 2, (n * nkb + i) / t32s_kps<P>());   // series 2: stage seen full by the MMA thread (uniform nkb only)
        if (leader) {
#pragma unroll
        for (int sub = 0; sub < t32s_kps<P>(); ++sub) {
          const uint32_t a_lo = (uint32_t)a_d0 + s * (STAGE >> 4) + sub * (A_TILE >> 4);
          const uint32_t b_lo = RKB > 0 ? b_lo_res + sub * (B_TILE >> 4) : (uint32_t)b_d0 + s * (STAGE >> 4) + sub * (B_TILE >> 4);
#pragma unroll
          for (int ks = 0; ks < kT32BK / 8; ++ks) {
            if constexpr (t32_f16<P>::value)
              t32_mma_f16(d_tmem, ((uint64_t)a_hi << 32) | (a_lo + ks * A_KS), ((uint64_t)b_hi << 32) | (b_lo + ks * B_KS), idesc,
                          i > 0 || sub > 0 || ks > 0);
            else
              t32_mma(d_tmem, ((uint64_t)a_hi << 32) | (a_lo + ks * A_KS), ((uint64_t)b_hi << 32) | (b_lo + ks * B_KS), idesc,
                      i > 0 || sub > 0 || ks > 0);
          }
        }
        t32_commit_a(empty_bar + 8 * s);   // slot free once these MMAs have read it
        }
        b_lo_res += t32s_kps<P>() * (B_TILE >> 4);
        if (++s == (uint32_t)S) { s = 0; ph ^= 1; }
      }
      if (leader) t32_commit_a(acc_full + 8 * a);
      t32s_trace(sk.dbg, 3, n);   // series 3: tile n fully issued
    }
  } else if (warp >= kT32SEpiWarp && warp < kT32SEpiWarp + 4) {
    // ------------------------------------------------------------------ epilogue
    const int quarter = warp & 3, r = quarter * 32 + (t & 31);   // accumulator row = TMEM lane
    uint32_t n = 0;
    for (int item = blockIdx.x; item < w.items; item += gridDim.x, ++n) {
      const T32Item it = t32s_item<BN>(w, item);
      const uint32_t a = n & 1;
      mbar_wait_a(acc_full + 8 * a, (n >> 1) & 1);
      t32_fence_after();
      if (t == kT32SEpiWarp * 32) t32s_trace(sk.dbg, 4, n);   // series 4: epilogue sees tile n complete
      const uint32_t lane_addr = tmem + a * COLS + ((uint32_t)(quarter * 32) << 16);
      const int m = it.m0 + r;
#pragma unroll 1
      for (int c = 0; c < BN; c += 16) {
        float v[16];
        t32_ld16(lane_addr + c, v);
        if (m < p.M) {
#pragma unroll
          for (int i = 0; i < 16; i += 4)
            if (it.n0 + c + i < p.N && !T32S_KNOB(w, 1)) p.store4(it.z, it.split, m, it.n0 + c + i, make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]));
        }
      }
      t32_fence_before();
      mbar_arrive_warp_a(acc_empty + 8 * a);
      if (t == kT32SEpiWarp * 32) t32s_trace(sk.dbg, 5, n);   // series 5: epilogue done with tile n
    }
  }

  t32_fence_before();
  __syncthreads();
  if (warp == kT32SMmaWarp) t32_dealloc<2 * COLS>(tmem);
  tl_end(sk.tl, tl_t0);
}

template <class P>
inline int t32s_round_kb(int kb) { return (kb + t32s_kps<P>() - 1) / t32s_kps<P>() * t32s_kps<P>(); }

// Split-K only for DEFER_REDUCE problems, and only as far as it takes to give every SM an item.
template <class P>
inline int t32s_splits_for(const P& p, int Z) {
  if (!P::DEFER_REDUCE) return 1;
  const int kblocks = t32s_round_kb<P>((p.K + kT32BK - 1) / kT32BK);
  const int tiles = ((p.M + kT32BM - 1) / kT32BM) * ((p.N + P::BN - 1) / P::BN) * Z;
  int s = kNumSMs / (tiles > 0 ? tiles : 1);
  const int cap = kblocks / 8 > 1 ? kblocks / 8 : 1;   // at least 8 k-blocks per split
  s = s < 1 ? 1 : (s > cap ? cap : s);
  const int per = t32s_round_kb<P>((kblocks + s - 1) / s);
  return (kblocks + per - 1) / per;
}

template <class P>
inline cudaError_t launch_gemm_tf32_stream(const P& p, int Z, int splits, SplitK sk, cudaStream_t st, int priority = 0) {
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(gemm_tf32_stream<P>, cudaFuncAttributeMaxDynamicSharedMemorySize, t32s_smem_bytes<P>());
    if (e != cudaSuccess) {
      cudaGetLastError();
      return e;
    }
    configured = true;
  }
  static_assert(t32s_smem_bytes<P>() <= 227 * 1024, "ring does not fit in shared memory");
  static_assert(t32s_stages<P>() >= 2, "ring too short");
  static_assert(t32_b_resident<P>::value % t32s_kps<P>() == 0, "resident B must hold whole stages");
  if (splits < 1 || (splits > 1 && !P::DEFER_REDUCE)) return cudaErrorInvalidValue;
  if (t32_b_resident<P>::value > 0 &&
      (splits != 1 || p.N > P::BN || (p.K + kT32BK - 1) / kT32BK > t32_b_resident<P>::value))
    return cudaErrorInvalidValue;   // resident B: one N tile, one split, K within the resident window
  T32Work w;
  w.tiles_m = (p.M + kT32BM - 1) / kT32BM;
  w.tiles_n = (p.N + P::BN - 1) / P::BN;
  w.kblocks = t32s_round_kb<P>((p.K + kT32BK - 1) / kT32BK);   // whole stages; loaders zero-fill k >= K
  w.per = t32s_round_kb<P>((w.kblocks + splits - 1) / splits);
  if (w.kblocks < 1 || (w.kblocks + w.per - 1) / w.per != splits) return cudaErrorInvalidValue;   // no empty splits
  w.splits = splits;
  w.fd_splits = t32s_fastdiv((uint32_t)splits); w.fd_tn = t32s_fastdiv((uint32_t)w.tiles_n); w.fd_tm = t32s_fastdiv((uint32_t)w.tiles_m);
  const long long items = (long long)w.tiles_m * w.tiles_n * splits * Z;
  if (items < 1 || items > 0x7fffffffLL) return cudaErrorInvalidValue;
  w.items = (int)items;
  { const char* e = getenv("DQNB_STREAM_DBG"); w.dbg = e ? atoi(e) : 0; }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(w.items < kNumSMs ? w.items : kNumSMs));
  cfg.blockDim = dim3(kT32SThreads);
  cfg.dynamicSmemBytes = t32s_smem_bytes<P>();
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributePriority;
  attr[0].val.priority = priority;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, gemm_tf32_stream<P>, p, sk, w);
}

}  // namespace dqnb
