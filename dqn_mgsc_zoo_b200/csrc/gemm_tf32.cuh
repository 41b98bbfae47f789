// tcgen05 TF32 tile GEMM for the latency-bound learner step ("slab" kernel).
//
// C[M,N] (+)= A[M,K] * B[N,K]^T with fp32 data fed to the 5th-gen tensor cores as TF32
// (tcgen05.mma.kind::tf32, fp32 accumulators in TMEM).  One CTA computes one 128 x BN tile
// over its split-K range of at most KB_MAX 32-float k-blocks.  Because that range is short,
// there is no smem ring: warps 0-7 issue *all* operand copies of the CTA up front (16-byte
// cp.async straight into the 128-byte-swizzled canonical UMMA layouts; uint8 sources go
// through registers), one commit group per k-block; each thread then retires its groups in
// order, proxy-fences and arrives on that k-block's mbarrier.  Warp 8's elected lane waits
// per k-block and issues the MMAs, so tensor work overlaps the tail of the loads; a final
// tcgen05.commit releases the epilogue, where thread t < 128 owns accumulator row t (TMEM lane t).
// Every barrier is used exactly once (parity 0): no phase bookkeeping.
//
// A problem type P provides
//   static constexpr int BN, KB_MAX;                BN in {32, 64, 128}
//   static constexpr bool A_MN_MAJOR, B_MN_MAJOR;   operand source contiguous along M/N instead of K
//   static constexpr bool DEFER_REDUCE;             split-K partials are left for a later kernel
//   int M, N, K;
//   ARow a_row(int z, int i) const;   decoded context of row i (K-major) / rows i..i+3 (MN-major)
//   ACol a_col(int z, int j) const;   decoded context of k = j..j+3 (K-major) / k = j (MN-major)
//   void a_put(const ARow&, const ACol&, uint32_t dst) const;   the 16-byte chunk -> smem (zero-fill OOB)
//   BRow b_row / BCol b_col / b_put: same for B[n, k]
//   void store4(int z, int split, int m, int n, float4 v) const;   C[m, n..n+3]: the finished sum, or this
//                                                                   split's partial when DEFER_REDUCE
// grid = (tiles_m * tiles_n, splits, Z), block = 288 threads.
#pragma once
#include <type_traits>
#include <utility>

#include "gemm_simt.cuh"

namespace dqnb {

constexpr int kT32Producers = 256; // warps 0-7 stage operands; warps 0-3 also run the epilogue
constexpr int kT32Threads = 288;   // + warp 8, whose elected lane issues the MMAs
constexpr int kT32BM = 128;
constexpr int kT32BK = 32;         // fp32 per k-block = one 128-byte swizzle row

// ---- staging primitives (dst = shared-space byte address of a 16-byte chunk) ----
__device__ __forceinline__ void cp16(uint32_t dst, const void* src, bool valid) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(valid ? 16 : 0) : "memory");
}
__device__ __forceinline__ void st16(uint32_t dst, float a, float b, float c, float d) {
  asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(dst), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ void cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// ---- descriptors (bit layouts: cute/arch/mma_sm100_desc.hpp) ----
// layout_type: 2 = SWIZZLE_128B (K-major operands), 1 = SWIZZLE_128B_BASE32B - the only smem
// layout tcgen05 accepts for MN-major 32-bit operands (cutlass sm100_common.inl:92).
__device__ __forceinline__ uint64_t t32_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes,
                                             uint32_t layout_type = 2) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;    // descriptor version (Blackwell)
  d |= (uint64_t)layout_type << 61;
  return d;
}
// kind::tf32: D=f32 (1<<4), A=B=TF32 (2<<7, 2<<10), majors, N>>3, M>>4.
__device__ __forceinline__ uint32_t t32_idesc(int n, bool a_mn, bool b_mn) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((a_mn ? 1u : 0u) << 15) | ((b_mn ? 1u : 0u) << 16) |
         ((uint32_t)(n >> 3) << 17) | ((uint32_t)(kT32BM >> 4) << 24);
}
__device__ __forceinline__ void t32_mma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, bool acc) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"((uint32_t)acc)
      : "memory");
}
// kind::f16 twin (fp16 operands, fp32 accumulate): K = 16 per instruction, i.e. the same 32 bytes of a
// K-major SWIZZLE_128B row as kind::tf32's K = 8, so descriptors and their k-step are shared.
__device__ __forceinline__ uint32_t t32_idesc_f16(int n, bool b_mn = false) {
  return (1u << 4) | ((b_mn ? 1u : 0u) << 16) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(kT32BM >> 4) << 24);   // D = f32, A = B = F16
}
__device__ __forceinline__ void t32_mma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, bool acc) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"((uint32_t)acc)
      : "memory");
}
__device__ __forceinline__ void t32_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void t32_commit_a(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void t32_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void t32_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
template <int kCols>
__device__ __forceinline__ void t32_alloc(uint32_t* smem_slot) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)), "n"(kCols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int kCols>
__device__ __forceinline__ void t32_dealloc(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols) : "memory");
}
// 16 consecutive fp32 columns of this thread's TMEM lane.
__device__ __forceinline__ void t32_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// ---- thread-block cluster primitives (split-K partials are reduced through distributed smem)
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_map(uint32_t local_smem_addr, uint32_t cta_rank) {
  uint32_t remote;
  asm("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_smem_addr), "r"(cta_rank));
  return remote;
}
// volatile (ordered after the barrier asm) but no memory clobber: the data is stable between the
// two cluster barriers, so the compiler may still batch the epilogue's global loads around it.
__device__ __forceinline__ float4 ld_cluster_f4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared::cluster.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ float4 ld_shared_f4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}
// this CTA's own shared::cta address as a shared::cluster address
__device__ __forceinline__ uint32_t local_as_cluster(uint32_t local_smem_addr) {
  uint32_t rank;
  asm("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  uint32_t remote;
  asm("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_smem_addr), "r"(rank));
  return remote;
}
constexpr int kT32MaxCluster = 8;   // portable cluster size limit = most split-K CTAs per tile

__host__ __device__ constexpr int t32_tmem_cols(int n) { return n <= 32 ? 32 : n <= 64 ? 64 : n <= 128 ? 128 : 256; }
__host__ __device__ constexpr int t32_tile_bytes(int rows, bool mn_major) {
  // K-major: [rows][128 B].  MN-major: [ceil(rows/32) mn-blocks][8 k-groups][4 k][128 B]
  return mn_major ? ((rows + 31) / 32) * 4096 : rows * 128;
}
template <class P>
__host__ __device__ constexpr int t32_smem_bytes() {
  return P::KB_MAX * (t32_tile_bytes(kT32BM, P::A_MN_MAJOR) + t32_tile_bytes(P::BN, P::B_MN_MAJOR)) + 1024;
}

// Which 16-byte chunks of a [ROWS x 32] operand k-block producer thread t stages.
//   K-major : the 16-byte column c = t%8 of rows t/8 + 32u           (u < ROWS/32)
//   MN-major: the mn-chunk cc = t % CH of k rows t/CH + (256/CH) u   (CH = ROWS/4, u < CH/8)
// so a thread meets only NROW distinct rows and NCOL distinct k positions per k-block: the
// problem's index decoding (a_row / a_col) is done once per context, not once per chunk, and
// consecutive threads read consecutive chunks of the contiguous source dimension.
template <int ROWS, bool MN_MAJOR>
struct T32Map {
  static_assert(ROWS % 32 == 0 && ROWS <= 128, "operand tile rows must be 32, 64, 96 or 128");
  static constexpr int CH = ROWS / 4;
  static constexpr int NROW = MN_MAJOR ? 1 : ROWS / 32;
  static constexpr int NCOL = MN_MAJOR ? CH / 8 : 1;
  static_assert(!MN_MAJOR || (kT32Producers % CH == 0 && CH % 8 == 0), "MN-major tile rows must be 32, 64 or 128");
  __device__ static __forceinline__ int row_index(int t, int u) { return MN_MAJOR ? 4 * (t % CH) : (t >> 3) + 32 * u; }
  __device__ static __forceinline__ int col_index(int t, int u) {
    return MN_MAJOR ? t / CH + (kT32Producers / CH) * u : 4 * (t & 7);
  }
  __device__ static __forceinline__ uint32_t offset(int t, int ur, int uc) {
    if (!MN_MAJOR) {
      int row = (t >> 3) + 32 * ur, c = t & 7;
      return (row >> 3) * 1024 + (row & 7) * 128 + ((c ^ (row & 7)) << 4);
    } else {
      // SWIZZLE_128B_BASE32B atom = 4 k-rows x 128 B, 32-byte units XOR-ed with the k-row
      int cc = t % CH, kk = t / CH + (kT32Producers / CH) * uc;
      int j = cc >> 3, c32 = (cc >> 1) & 3, h = cc & 1, g = kk >> 2, kr = kk & 3;
      return j * 4096 + g * 512 + kr * 128 + ((c32 ^ kr) << 5) + (h << 4);
    }
  }
};

// Optional members of a problem type, detected here:
//   static constexpr bool A_REG = true;  uint32_t a_get(row, col) const;  void a_store(dst, word) const;
// A operand chunks come through registers (4 packed uint8 -> 4 floats).  All loads of the CTA's
// K range are issued before the first use so they share one memory round trip.
template <class P, class = void> struct t32_a_reg : std::false_type {};
template <class P> struct t32_a_reg<P, std::enable_if_t<P::A_REG>> : std::true_type {};
//   static constexpr bool A_ROW16 = true;  const uint8_t* a_row16(z, m) const;  int a_kb_off(kb) const;
// K-major uint8 A whose k-block of a row is 32 contiguous source bytes (conv1: one kernel row of the
// patch = 8 px x 4 ch): thread t owns half a row (16 bytes = one LDG.128) of every k-block, i.e.
// 4x fewer loads than the word-per-chunk path; the 4 widened chunks go to swizzled smem.
template <class P, class = void> struct t32_a_row16 : std::false_type {};
//   static constexpr bool F16 = true;  both operands are fp16, K-major; a k-block is 64 halves (128 bytes per
//   row, like 32 floats) and p.K counts 32-bit words.  With A_ROW16H the uint8 A rows (two 32-byte patch
//   rows per k-block) are widened to halves: thread t owns bytes [16 half, 16 half + 16) of both.
template <class P, class = void> struct t32_f16 : std::false_type {};
template <class P> struct t32_f16<P, std::enable_if_t<P::F16>> : std::true_type {};
template <class P, class = void> struct t32_a_row16h : std::false_type {};
template <class P> struct t32_a_row16h<P, std::enable_if_t<P::A_ROW16H>> : std::true_type {};
// 8 packed uint8 (two words) -> 8 halves without conversions: PRMT drops a byte into the mantissa of 1024
// (0x6400 | b is 1024 + b exactly), HSUB2 removes the 1024.
__device__ __forceinline__ uint32_t t32_u8x2_to_half2(uint32_t w, uint32_t selector) {
  uint32_t r;
  asm("{\n.reg .b32 t, c;\nmov.b32 c, 0x64006400;\nprmt.b32 t, %1, 0x64646464, %2;\nsub.f16x2 %0, t, c;\n}\n" : "=r"(r) : "r"(w), "r"(selector));
  return r;
}
__device__ __forceinline__ void t32_widen8h(uint32_t dst, uint32_t w0, uint32_t w1) {
  asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dst), "r"(t32_u8x2_to_half2(w0, 0x4140)),
               "r"(t32_u8x2_to_half2(w0, 0x4342)), "r"(t32_u8x2_to_half2(w1, 0x4140)), "r"(t32_u8x2_to_half2(w1, 0x4342))
               : "memory");
}
template <class P> struct t32_a_row16<P, std::enable_if_t<P::A_ROW16>> : std::true_type {};
//   static constexpr bool PRELOAD = true;
//   Ctx  store_ctx(z, split, n) const;                      per thread: what depends on the column only (destination bases, bias)
//   Pre  preload(const Ctx&, z, split, m, n) const;         per row: element offsets + the global loads the store needs
//   void store4p(const Ctx&, const Pre&, float4 v) const;   the store itself: arithmetic and st.global only
// Both are evaluated while the MMAs are still running, so that neither the round trip of those loads (1 000-2 000
// clocks from L2) nor the address arithmetic (~60 instructions per row when left to store4) sits between the finished
// accumulators and the stores.  store4 stays: the stream kernel and the non-PRELOAD instantiations call it.
struct T32NoPreload { struct Pre {}; struct Ctx {}; };
template <class P, class = void> struct t32_preload : std::false_type {};
template <class P> struct t32_preload<P, std::enable_if_t<P::PRELOAD>> : std::true_type {};
//   static constexpr bool B_TMA = true;  const void* b_tmap(int z) const;
// B is a plain fp32 matrix described by a tiled tensor map (K-major: dims {K, N}, box {32, BN}, SWIZZLE_128B;
// MN-major: dims {N, K}, box {32, 32}, SWIZZLE_128B_ATOM_32B - the two canonical UMMA layouts above are exactly
// what the TMA engine writes in those modes).  The MMA issuer thread then fetches every B k-block of the CTA with
// cp.async.bulk.tensor onto that k-block's barrier and the eight producer warps stage only A.  A null map (weights
// that are not one of the learner's own shadows) falls back to the cp.async path at run time.
template <class P, class = void> struct t32_b_tma : std::false_type {};
template <class P> struct t32_b_tma<P, std::enable_if_t<P::B_TMA>> : std::true_type {};

// Retire this thread's commit groups oldest-first: when group I has landed, make it visible to
// the tensor core (async proxy) and arrive on k-block I's barrier.
template <int KB, int I = 0>
__device__ __forceinline__ void t32_retire_groups(int nkb, uint64_t* full_bar) {
  if constexpr (I < KB) {
    if (I < nkb) {
      cp_wait<KB - 1 - I>();
      fence_proxy_async_smem();
      mbar_arrive_warp(&full_bar[I]);
    }
    t32_retire_groups<KB, I + 1>(nkb, full_bar);
  }
}

// The same for ONE k-block whose index is a compile-time constant (register-staged A operands: the thread has
// just written its chunks of k-block I with st.shared; B's group I is waited for, then both become visible).
template <int KB, int I>
__device__ __forceinline__ void t32_retire_one(int nkb, uint64_t* full_bar) {
  if (I < nkb) {
    cp_wait<KB - 1 - I>();
    fence_proxy_async_smem();
    mbar_arrive_warp(&full_bar[I]);
  }
}
// cp.async pipeline of the generic producer path: k-block i is published kT32Lag iterations after it was issued
// (its copies have had that long to land), so the MMA issuer works on the first k-blocks while the producers are
// still issuing the later ones - before, every arrival came after the whole slab had been issued, which put the
// CTA's 30-odd MMAs (~96 clocks each) behind the issue phase instead of under it.
constexpr int kT32Lag = 3;
template <int J = kT32Lag - 1>
__device__ __forceinline__ void t32_drain_groups(int nkb, uint64_t* full_bar) {
  if constexpr (J >= 0) {
    if (nkb - 1 - J >= 0) {
      cp_wait<J>();
      fence_proxy_async_smem();
      mbar_arrive_warp(&full_bar[nkb - 1 - J]);
    }
    t32_drain_groups<J - 1>(nkb, full_bar);
  }
}

//   static constexpr bool EARLY_TRIGGER = true;   griddepcontrol.launch_dependents at the START of the kernel
// (forward chain, where no side-branch kernel competes for the idle SMs): the next kernel's CTAs are placed on
// whatever SMs this grid leaves free, set up their barriers / TMEM and pull their weight operands by TMA while this
// kernel still computes; they block at their own griddepcontrol.wait until this grid has completed.
template <class P, class = void> struct t32_early_trigger : std::false_type {};
template <class P> struct t32_early_trigger<P, std::enable_if_t<P::EARLY_TRIGGER>> : std::true_type {};
//   static constexpr int MIN_CTAS = n;   register budget for n resident CTAs per SM (__launch_bounds__)
template <class P, class = void> struct t32_min_ctas : std::integral_constant<int, 1> {};
template <class P> struct t32_min_ctas<P, std::enable_if_t<(P::MIN_CTAS > 1)>> : std::integral_constant<int, P::MIN_CTAS> {};

template <class P>
__global__ void __launch_bounds__(kT32Threads, t32_min_ctas<P>::value) gemm_tf32(P p, SplitK sk) {
  constexpr int BN = P::BN, COLS = t32_tmem_cols(BN), KB = P::KB_MAX;
  constexpr int A_TILE = t32_tile_bytes(kT32BM, P::A_MN_MAJOR), B_TILE = t32_tile_bytes(BN, P::B_MN_MAJOR);
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t full_bar[KB];
  __shared__ uint64_t done_bar;
  __shared__ uint32_t tmem_slot;

  const int t = threadIdx.x, warp = t >> 5;
  const unsigned long long tl_t0 = tl_begin(sk.tl);
  const bool dbg = sk.dbg != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0;
  if (dbg && t == 0) sk.dbg[0] = clock64();
  const uint32_t smem = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int tiles_n = (p.N + BN - 1) / BN;
  const int tile = blockIdx.x, tile_m = tile / tiles_n, tile_n = tile - tile_m * tiles_n;
  const int z = blockIdx.z, splits = gridDim.y, split = blockIdx.y;
  const int m0 = tile_m * kT32BM, n0 = tile_n * BN;
  const int kblocks = (p.K + kT32BK - 1) / kT32BK;
  const int per = (kblocks + splits - 1) / splits;      // host guarantees per <= KB
  const int kb_beg = split * per, kb_end = min(kblocks, kb_beg + per);
  const int nkb = max(kb_end - kb_beg, 0);

  const void* bmap = nullptr;
  if constexpr (t32_b_tma<P>::value) bmap = p.b_tmap(z);
  const bool b_tma = bmap != nullptr;   // B arrives by TMA: one more arrival (the expect_tx) per k-block
  if (t == 0) {
    for (int s = 0; s < KB; ++s) mbar_init(&full_bar[s], kT32Producers / 32 + (b_tma ? 1 : 0));
    mbar_init(&done_bar, 1);
    fence_mbar_init();
  }
  if (warp == 8) t32_alloc<COLS>(&tmem_slot);
  if (b_tma && warp == 8 && elect_one()) tma_prefetch_map(bmap);   // the descriptor itself is never written by a kernel
  t32_fence_before();
  __syncthreads();
  t32_fence_after();
  const uint32_t tmem = tmem_slot;
  // B by tensor-map TMA = the learner's own weight shadows, final before the update's graph started (the fused fc1
  // optimiser step that rewrites them is ordered behind every reader by events): every B k-block of this CTA is
  // requested NOW, ahead of griddepcontrol.wait - under programmatic dependent launch that is while the preceding
  // kernel is still storing, so a third to a half of the CTA's operand bytes arrive before its inputs even exist.
  if constexpr (t32_b_tma<P>::value) {
    if (b_tma && warp == 8 && elect_one()) {
      for (int i = 0; i < nkb; ++i) {
        const int k0 = (kb_beg + i) * (t32_f16<P>::value ? 2 * kT32BK : kT32BK);   // in elements of the map's type
        const uint32_t tile = smem + KB * A_TILE + i * B_TILE, bar = smem_u32(&full_bar[i]);
        mbar_expect_tx(&full_bar[i], B_TILE);
        if constexpr (P::B_MN_MAJOR && t32_f16<P>::value) {   // 16-bit MN-major: atoms of 64 n x 64 k
#pragma unroll
          for (int j = 0; j < BN / 64; ++j) tma_load_2d(tile + j * 8192, bmap, n0 + 64 * j, k0, bar);
        } else if constexpr (P::B_MN_MAJOR) {
#pragma unroll
          for (int j = 0; j < BN / 32; ++j) tma_load_2d(tile + j * 4096, bmap, n0 + 32 * j, k0, bar);
        } else {
          tma_load_2d(tile, bmap, k0, n0, bar);
        }
      }
    }
  }
  // everything above touched only this CTA's shared memory / TMEM and constant weights; from here on the kernel
  // reads what its predecessors wrote
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if constexpr (t32_early_trigger<P>::value) asm volatile("griddepcontrol.launch_dependents;");
  if (dbg && t == 0) sk.dbg[1] = clock64();     // setup done

  if (warp < 8) {
    // ---- producers: everything in flight at once, one commit group per k-block
    using MA = T32Map<kT32BM, P::A_MN_MAJOR>;
    using MB = T32Map<BN, P::B_MN_MAJOR>;
    typename P::ARow arow[MA::NROW];
    typename P::BRow brow[MB::NROW];
#pragma unroll
    for (int u = 0; u < MA::NROW; ++u) arow[u] = p.a_row(z, m0 + MA::row_index(t, u));
#pragma unroll
    for (int u = 0; u < MB::NROW; ++u) brow[u] = p.b_row(z, n0 + MB::row_index(t, u));
    auto stage_b = [&](int i) {
      const int k0 = (kb_beg + i) * kT32BK;
      const uint32_t tile = smem + KB * A_TILE + i * B_TILE;
#pragma unroll
      for (int uc = 0; uc < MB::NCOL; ++uc) {
        typename P::BCol col = p.b_col(z, k0 + MB::col_index(t, uc));
#pragma unroll
        for (int ur = 0; ur < MB::NROW; ++ur) p.b_put(brow[ur], col, tile + MB::offset(t, ur, uc));
      }
    };
    if constexpr (t32_a_row16h<P>::value) {
      static_assert(!P::A_MN_MAJOR && t32_f16<P>::value, "A_ROW16H is the K-major fp16 path");
      const int row = t >> 1, half = t & 1;
      const uint8_t* rp = p.a_row16(z, m0 + row);
      uint4 w[KB][2];
#pragma unroll
      for (int i = 0; i < KB; ++i) {
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          w[i][s] = make_uint4(0u, 0u, 0u, 0u);
          if (i < nkb && rp != nullptr)
            w[i][s] = __ldg(reinterpret_cast<const uint4*>(rp + p.a_kb_off(2 * (kb_beg + i) + s) + half * 16));
        }
      }
#pragma unroll 1
      for (int i = 0; i < nkb; ++i) {
        if (!b_tma) stage_b(i);
        cp_commit();
      }
#pragma unroll 1
      for (int i = nkb; i < KB; ++i) cp_commit();     // empty groups keep the wait counts static
      const uint32_t row_off = (row >> 3) * 1024 + (row & 7) * 128;
      [&]<int... I>(std::integer_sequence<int, I...>) {
        ([&] {
          if (I < nkb) {
            const uint32_t base = smem + I * A_TILE + row_off;
#pragma unroll
            for (int s = 0; s < 2; ++s) {
              const uint32_t c = s * 4 + half * 2;   // 16-byte chunk (8 halves) inside the 128-byte row
              t32_widen8h(base + ((c ^ (uint32_t)(row & 7)) << 4), w[I][s].x, w[I][s].y);
              t32_widen8h(base + (((c + 1) ^ (uint32_t)(row & 7)) << 4), w[I][s].z, w[I][s].w);
            }
          }
          t32_retire_one<KB, I>(nkb, full_bar);     // k-block I is complete: the MMAs on it start now
        }(), ...);
      }(std::make_integer_sequence<int, KB>{});
    } else if constexpr (t32_a_row16<P>::value) {
      static_assert(!P::A_MN_MAJOR, "A_ROW16 is a K-major path");
      const int row = t >> 1, half = t & 1;
      const uint8_t* rp = p.a_row16(z, m0 + row);
      uint4 w[KB];
#pragma unroll
      for (int i = 0; i < KB; ++i) {
        w[i] = make_uint4(0u, 0u, 0u, 0u);
        if (i < nkb && rp != nullptr) w[i] = __ldg(reinterpret_cast<const uint4*>(rp + p.a_kb_off(kb_beg + i) + half * 16));
      }
#pragma unroll 1
      for (int i = 0; i < nkb; ++i) {
        if (!b_tma) stage_b(i);
        cp_commit();
      }
#pragma unroll 1
      for (int i = nkb; i < KB; ++i) cp_commit();     // empty groups keep the wait counts static
      const uint32_t row_off = (row >> 3) * 1024 + (row & 7) * 128;
      [&]<int... I>(std::integer_sequence<int, I...>) {
        ([&] {
          if (I < nkb) {
            const uint32_t base = smem + I * A_TILE + row_off;
            const uint32_t wd[4] = {w[I].x, w[I].y, w[I].z, w[I].w};
#pragma unroll
            for (int j = 0; j < 4; ++j) p.a_store(base + ((uint32_t)((half * 4 + j) ^ (row & 7)) << 4), wd[j]);
          }
          t32_retire_one<KB, I>(nkb, full_bar);
        }(), ...);
      }(std::make_integer_sequence<int, KB>{});
    } else if constexpr (t32_a_reg<P>::value) {
      constexpr int NW = MA::NROW * MA::NCOL;
      uint32_t word[KB * NW];
#pragma unroll
      for (int i = 0; i < KB; ++i) {
#pragma unroll
        for (int uc = 0; uc < MA::NCOL; ++uc) {
          typename P::ACol col = p.a_col(z, (kb_beg + i) * kT32BK + MA::col_index(t, uc));
#pragma unroll
          for (int ur = 0; ur < MA::NROW; ++ur)
            word[(i * MA::NCOL + uc) * MA::NROW + ur] = i < nkb ? p.a_get(arow[ur], col) : 0u;
        }
      }
#pragma unroll 1
      for (int i = 0; i < nkb; ++i) {
        if (!b_tma) stage_b(i);
        cp_commit();
      }
#pragma unroll 1
      for (int i = nkb; i < KB; ++i) cp_commit();     // empty groups keep the wait counts static
      [&]<int... I>(std::integer_sequence<int, I...>) {
        ([&] {
#pragma unroll
          for (int uc = 0; uc < MA::NCOL; ++uc)
#pragma unroll
            for (int ur = 0; ur < MA::NROW; ++ur)
              if (I < nkb) p.a_store(smem + I * A_TILE + MA::offset(t, ur, uc), word[(I * MA::NCOL + uc) * MA::NROW + ur]);
          t32_retire_one<KB, I>(nkb, full_bar);
        }(), ...);
      }(std::make_integer_sequence<int, KB>{});
    } else {
#pragma unroll 1
      for (int i = 0; i < nkb; ++i) {
        const int k0 = (kb_beg + i) * kT32BK;
        const uint32_t tile = smem + i * A_TILE;
#pragma unroll
        for (int uc = 0; uc < MA::NCOL; ++uc) {
          typename P::ACol col = p.a_col(z, k0 + MA::col_index(t, uc));
#pragma unroll
          for (int ur = 0; ur < MA::NROW; ++ur) p.a_put(arow[ur], col, tile + MA::offset(t, ur, uc));
        }
        if (!b_tma) stage_b(i);
        cp_commit();
        if (i >= kT32Lag) {          // publish k-block i - kT32Lag: at most the kT32Lag newest groups may still be in flight
          cp_wait<kT32Lag>();
          fence_proxy_async_smem();
          mbar_arrive_warp(&full_bar[i - kT32Lag]);
        }
      }
      if (dbg && t == 0) sk.dbg[2] = clock64();     // all copies issued
      t32_drain_groups<>(nkb, full_bar);
    }
    if constexpr (t32_a_row16h<P>::value || t32_a_row16<P>::value || t32_a_reg<P>::value) {
      if (dbg && t == 0) sk.dbg[2] = clock64();     // (register paths: issued and landed coincide)
    }
    if (dbg && t == 0) sk.dbg[3] = clock64();     // this thread's copies landed
  } else if (warp == 8) {
    // ---- MMA issuer: the whole warp walks the barriers and one ELECTED lane issues - with descriptors formed as a
    // warp-uniform base plus a compile-time offset ptxas keeps them in uniform registers (a `threadIdx.x == 256`
    // test made it move every operand through a R2UR.BROADCAST waterfall: 18 instructions per tcgen05.mma)
    const bool leader = elect_one();
    constexpr bool F16 = t32_f16<P>::value;
    static_assert(!(F16 && P::A_MN_MAJOR), "fp16 A operands are K-major");
    const uint32_t idesc = F16 ? t32_idesc_f16(BN, P::B_MN_MAJOR) : t32_idesc(BN, P::A_MN_MAJOR, P::B_MN_MAJOR);
    const uint64_t a_d0 = P::A_MN_MAJOR ? t32_desc(smem, 4096, 512, 1) : t32_desc(smem, 16, 1024);
    // MN-major B: fp32 in SWIZZLE_128B_BASE32B atoms of 32 n x 4 k, fp16 in SWIZZLE_128B atoms of 64 n x 8 k
    const uint64_t b_d0 = !P::B_MN_MAJOR ? t32_desc(smem + KB * A_TILE, 16, 1024)
                          : F16 ? t32_desc(smem + KB * A_TILE, 8192, 1024, 2) : t32_desc(smem + KB * A_TILE, 4096, 512, 1);
    constexpr uint32_t A_KS = (P::A_MN_MAJOR ? 1024 : 32) >> 4, B_KS = (!P::B_MN_MAJOR ? 32 : F16 ? 2048 : 1024) >> 4;
#pragma unroll
    for (int i = 0; i < KB; ++i) {
      if (i < nkb) {
        mbar_wait(&full_bar[i], 0);
        t32_fence_after();
        if (leader) {
#pragma unroll
          for (int ks = 0; ks < kT32BK / 8; ++ks) {
            const uint64_t ad = a_d0 + (uint32_t)((i * A_TILE) >> 4) + ks * A_KS;
            const uint64_t bd = b_d0 + (uint32_t)((i * B_TILE) >> 4) + ks * B_KS;
            if constexpr (t32_f16<P>::value) t32_mma_f16(tmem, ad, bd, idesc, i > 0 || ks > 0);
            else t32_mma(tmem, ad, bd, idesc, i > 0 || ks > 0);
          }
        }
      }
    }
    if (leader) t32_commit(&done_bar);
  }

  // ---- the store assignment, fixed before the accumulators exist: thread t < 256 owns one float4 column group
  // (c4 = t mod G4) of rows row_beg + t / G4 + k * RPP.  The column - hence the bias / mask column the problem's
  // store reads - is fixed per thread, the row advances by a constant and the passes are fully unrolled.  (A rolled
  // loop with per-item division and store4's address arithmetic ran ~400 clocks per item at 8 warps per SM.)
  // Split-K CTAs of a tile form a cluster along grid.y (unless the problem defers the reduction): CTA `split` then
  // owns rows [split*R, (split+1)*R) of the summed tile; otherwise the CTA owns all 128 rows of its own tile.
  constexpr int PSTRIDE = BN + 4;   // floats; +4 keeps the float4 row writes off one bank
  constexpr int G4 = BN / 4;
  constexpr int RPP = kT32Producers / G4;          // rows per pass
  constexpr int PASSES = kT32BM / RPP;
  const bool clustered = !P::DEFER_REDUCE && splits > 1;
  const int c4 = t % G4, r0 = t / G4;
  const int n = n0 + c4 * 4;
  const bool col_ok = n < p.N;
  const int rows_here = p.M - m0;                  // rows of this tile that exist
  const int R = clustered ? (kT32BM + splits - 1) / splits : kT32BM;
  const int row_beg = clustered ? split * R : 0;
  const int row_end = min(min(kT32BM, row_beg + R), rows_here);
  // What the problem's store reads from global memory (bias, ReLU mask of the layer input, optimiser state) is
  // fetched NOW, while the tensor core is still working: issued after the accumulators are ready, that round trip
  // (1 000-2 000 clocks from L2) was the longest part of the store phase.
  using PreOwner = std::conditional_t<t32_preload<P>::value, P, T32NoPreload>;
  typename PreOwner::Ctx sctx;
  typename PreOwner::Pre pre[PASSES];
  if constexpr (t32_preload<P>::value) {
    if (warp < 8 && col_ok) {
      sctx = p.store_ctx(z, split, n);
#pragma unroll
      for (int k = 0; k < PASSES; ++k) {
        const int r = row_beg + r0 + k * RPP;
        if (r < row_end) pre[k] = p.preload(sctx, z, split, m0 + r, n);
      }
    }
  }
  // ---- epilogue, part 1: accumulators TMEM -> registers -> this CTA's smem tile [128][BN+4]
  // (the operand slab is dead once the MMAs are done).  Thread t < 128 owns row t (TMEM lane t).
  if (warp < 4) {
    if (nkb > 0) mbar_wait(&done_bar, 0);
    t32_fence_after();
    if (dbg && t == 0) sk.dbg[4] = clock64();     // MMAs complete
    const uint32_t lane_addr = tmem + ((uint32_t)(warp * 32) << 16);
#pragma unroll 1
    for (int c = 0; c < BN; c += 16) {
      float v[16];
      if (nkb > 0) t32_ld16(lane_addr + c, v);
      else {
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = 0.f;
      }
#pragma unroll
      for (int i = 0; i < 16; i += 4) st16(smem + (t * PSTRIDE + c + i) * 4, v[i], v[i + 1], v[i + 2], v[i + 3]);
    }
  }
  if (dbg && t == 0) sk.dbg[5] = clock64();       // tile parked in smem
  // ---- part 2: the tile leaves as coalesced float4 items handed to the problem's store (bias / relu / mask /
  // partial-buffer placement live there); clustered tiles are summed in split order through distributed smem.
  if (clustered) cluster_sync_all(); else __syncthreads();
  // Programmatic dependent launch: the MMAs are done and only the store loop is left, so let the
  // next kernel of the stream get its CTAs placed and run its prologue (barriers, TMEM) now; it
  // blocks at its griddepcontrol.wait until this grid has completed.  Triggering earlier parks idle
  // CTAs on SMs the side-branch kernels need.  (No-ops for launches without the attribute.)
  if constexpr (!t32_early_trigger<P>::value) asm volatile("griddepcontrol.launch_dependents;");
  if (dbg && t == 0) sk.dbg[7] = clock64();
  if (warp < 8) {
    auto emit = [&](int k, int r, const float4& v) {
      if constexpr (t32_preload<P>::value) p.store4p(sctx, pre[k], v);
      else p.store4(z, split, m0 + r, n, v);
    };
    if (!clustered) {
      float4 v[PASSES];
#pragma unroll
      for (int k = 0; k < PASSES; ++k) v[k] = ld_shared_f4(smem + ((r0 + k * RPP) * PSTRIDE + c4 * 4) * 4);
#pragma unroll
      for (int k = 0; k < PASSES; ++k) {
        const int r = r0 + k * RPP;
        if (col_ok && r < row_end) emit(k, r, v[k]);
      }
    } else {
      const int nsum = splits;
#pragma unroll
      for (int k = 0; k < PASSES; ++k) {
        const int r = row_beg + r0 + k * RPP;
        if (r < row_end) {
          const uint32_t local = smem + (r * PSTRIDE + c4 * 4) * 4;
          // all partial tiles' loads in flight together (a DSMEM load is a few hundred cycles), then the
          // sum in split order
          float4 x[kT32MaxCluster];
#pragma unroll
          for (int q = 0; q < kT32MaxCluster; ++q)
            if (q < nsum) x[q] = ld_cluster_f4(cluster_map(local, (uint32_t)q));
          float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
          for (int q = 0; q < kT32MaxCluster; ++q)
            if (q < nsum) { v.x += x[q].x; v.y += x[q].y; v.z += x[q].z; v.w += x[q].w; }
          if (col_ok) emit(k, r, v);
        }
      }
    }
  }
  if (dbg && t == 0) sk.dbg[6] = clock64();       // my rows reduced and stored
  if (clustered) cluster_sync_all();   // nobody leaves (and frees its smem) while a peer may still read it

  t32_fence_before();
  __syncthreads();
  if (warp == 8) t32_dealloc<COLS>(tmem);
  tl_end(sk.tl, tl_t0);
}

// Split-K factor: at least what makes the K range fit the slab; more (down to 4 k-blocks per
// CTA) when the tile count alone would leave most of the 148 SMs idle.
template <class P>
inline int t32_splits_for(const P& p, int Z, int want_ctas = 120) {
  int kblocks = (p.K + kT32BK - 1) / kT32BK;
  int min_s = (kblocks + P::KB_MAX - 1) / P::KB_MAX;
  int tiles = ((p.M + kT32BM - 1) / kT32BM) * ((p.N + P::BN - 1) / P::BN) * Z;
  int s = want_ctas / (tiles > 0 ? tiles : 1);
  int cap = kblocks / 4 > min_s ? kblocks / 4 : min_s;
  s = s < min_s ? min_s : (s > cap ? cap : s);
  if (!P::DEFER_REDUCE && s > kT32MaxCluster) s = kT32MaxCluster;   // min_s > 8 is rejected at launch
  if (tiles * s > kNumSMs && kNumSMs / tiles >= min_s) s = kNumSMs / tiles;   // one CTA per SM: stay in one wave
  int per = (kblocks + s - 1) / s;
  return (kblocks + per - 1) / per;
}
template <class P>
// `priority`: CUDA launch priority (lower = scheduled first); kernels of the dependent chain get
// the device's greatest priority so that side-branch kernels do not take their SMs.
inline cudaError_t launch_gemm_tf32(const P& p, int Z, int splits, SplitK sk, cudaStream_t st, int priority = 0,
                                    bool pdl = false) {
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(gemm_tf32<P>, cudaFuncAttributeMaxDynamicSharedMemorySize, t32_smem_bytes<P>());
    if (e != cudaSuccess) {
      cudaGetLastError();   // do not leave the error for the next launch to trip over
      return e;
    }
    configured = true;
  }
  static_assert(t32_smem_bytes<P>() <= 227 * 1024, "slab does not fit in shared memory: lower KB_MAX or BN");
  static_assert(kT32BM * (P::BN + 4) * 4 <= t32_smem_bytes<P>() - 1024,
                "the accumulator tile must fit in the (dead) operand slab");
  int kblocks = (p.K + kT32BK - 1) / kT32BK;
  if ((kblocks + splits - 1) / splits > P::KB_MAX) return cudaErrorInvalidValue;
  const bool clustered = !P::DEFER_REDUCE && splits > 1;
  if (clustered && splits > kT32MaxCluster) return cudaErrorInvalidValue;
  int tiles = ((p.M + kT32BM - 1) / kT32BM) * ((p.N + P::BN - 1) / P::BN);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(tiles, splits, Z);
  cfg.blockDim = dim3(kT32Threads);
  cfg.dynamicSmemBytes = t32_smem_bytes<P>();
  cfg.stream = st;
  cudaLaunchAttribute attr[3];
  unsigned na = 0;
  if (pdl) {   // may start before the previous kernel of `st` has finished (see griddepcontrol.wait)
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na++].val.programmaticStreamSerializationAllowed = 1;
  }
  attr[na].id = cudaLaunchAttributePriority;
  attr[na++].val.priority = priority;
  if (clustered) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = 1; attr[na].val.clusterDim.y = splits; attr[na].val.clusterDim.z = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  return cudaLaunchKernelEx(&cfg, gemm_tf32<P>, p, sk);
}
template <class P>
inline size_t gemm_tf32_tiles(const P& p, int Z) {
  return (size_t)((p.M + kT32BM - 1) / kT32BM) * ((p.N + P::BN - 1) / P::BN) * Z;
}
template <class P>
inline size_t gemm_tf32_ws_floats(const P&, int, int) { return 0; }   // partials never touch global memory

}  // namespace dqnb
