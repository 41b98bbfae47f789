// tcgen05 TF32 tile GEMM for the latency-bound learner step ("slab" kernel).
//
// C[M,N] (+)= A[M,K] * B[N,K]^T with fp32 data fed to the 5th-gen tensor cores as TF32
// (tcgen05.mma.kind::tf32, fp32 accumulators in TMEM).  One CTA computes one 128 x BN tile
// over its split-K range of at most KB_MAX 32-float k-blocks.  Because that range is short,
// there is no smem ring: warps 0-3 issue *all* operand copies of the CTA up front (16-byte
// cp.async straight into the 128-byte-swizzled canonical UMMA layouts; uint8 sources are
// converted in registers), one commit group per k-block; each thread then retires its groups in
// order, proxy-fences and arrives on that k-block's mbarrier.  Warp 4's elected lane waits
// per k-block and issues the MMAs, so tensor work overlaps the tail of the loads; a final
// tcgen05.commit releases the epilogue, where thread t owns accumulator row t (TMEM lane t).
// Every barrier is used exactly once (parity 0): no phase bookkeeping.
//
// A problem type P provides
//   static constexpr int BN, KB_MAX;
//   static constexpr bool A_MN_MAJOR, B_MN_MAJOR;   operand source contiguous along M/N instead of K
//   static constexpr bool DEFER_REDUCE;             split-K partials are left for a later kernel
//   int M, N, K;
//   void a_chunk(int z, int i, int j, uint32_t dst) const;   K-major: A[m=i, k=j..j+3] -> 16 B at dst
//                                                            MN-major: A[m=i..i+3, k=j]     (zero-fill OOB)
//   void b_chunk(int z, int i, int j, uint32_t dst) const;   same for B[n, k]
//   void store_row(int z, int m, int n0, const float (&v)[16], int valid_n) const;
//   void store_partial(int z, int split, int m, int n0, const float (&v)[16], int valid_n) const;  (DEFER_REDUCE)
// grid = (tiles_m * tiles_n, splits, Z), block = 160 threads.
#pragma once
#include "gemm_simt.cuh"

namespace dqnb {

constexpr int kT32Threads = 160;   // 4 producer/epilogue warps + 1 MMA-issue warp
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
__device__ __forceinline__ void t32_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
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

__host__ __device__ constexpr int t32_tmem_cols(int n) { return n <= 32 ? 32 : n <= 64 ? 64 : n <= 128 ? 128 : 256; }
__host__ __device__ constexpr int t32_tile_bytes(int rows, bool mn_major) {
  // K-major: [rows][128 B].  MN-major: [ceil(rows/32) mn-blocks][8 k-groups][4 k][128 B]
  return mn_major ? ((rows + 31) / 32) * 4096 : rows * 128;
}
template <class P>
__host__ __device__ constexpr int t32_smem_bytes() {
  return P::KB_MAX * (t32_tile_bytes(kT32BM, P::A_MN_MAJOR) + t32_tile_bytes(P::BN, P::B_MN_MAJOR)) + 1024;
}

// Stage one [ROWS x 32] operand k-block.  K-major: chunk id -> (row = id/8, c = id%8);
// MN-major: (k = id / CH, mn-chunk = id % CH).  Consecutive threads take consecutive chunks
// of the contiguous source dimension.
template <int ROWS, bool MN_MAJOR, class F>
__device__ __forceinline__ void t32_stage(uint32_t tile, int r0, int k0, F&& chunk) {
  if (!MN_MAJOR) {
#pragma unroll 4
    for (int id = threadIdx.x; id < ROWS * 8; id += 128) {
      int row = id >> 3, c = id & 7;
      chunk(r0 + row, k0 + c * 4, tile + (row >> 3) * 1024 + (row & 7) * 128 + ((c ^ (row & 7)) << 4));
    }
  } else {
    constexpr int CH = ((ROWS + 31) / 32) * 8;   // 16-byte mn-chunks per k row
#pragma unroll 4
    for (int id = threadIdx.x; id < kT32BK * CH; id += 128) {
      int k = id / CH, cc = id - k * CH;
      // SWIZZLE_128B_BASE32B atom = 4 k-rows x 128 B, 32-byte units XOR-ed with the k-row
      int j = cc >> 3, c32 = (cc >> 1) & 3, h = cc & 1, g = k >> 2, kr = k & 3;
      chunk(r0 + cc * 4, k0 + k, tile + j * 4096 + g * 512 + kr * 128 + ((c32 ^ kr) << 5) + (h << 4));
    }
  }
}

// Retire this thread's commit groups oldest-first: when group I has landed, make it visible to
// the tensor core (async proxy) and arrive on k-block I's barrier.
template <int KB, int I = 0>
__device__ __forceinline__ void t32_retire_groups(int nkb, uint64_t* full_bar) {
  if constexpr (I < KB) {
    if (I < nkb) {
      cp_wait<KB - 1 - I>();
      fence_proxy_async_smem();
      mbar_arrive(&full_bar[I]);
    }
    t32_retire_groups<KB, I + 1>(nkb, full_bar);
  }
}

template <class P>
__global__ void __launch_bounds__(kT32Threads) gemm_tf32(P p, SplitK sk) {
  constexpr int BN = P::BN, COLS = t32_tmem_cols(BN), KB = P::KB_MAX;
  constexpr int A_TILE = t32_tile_bytes(kT32BM, P::A_MN_MAJOR), B_TILE = t32_tile_bytes(BN, P::B_MN_MAJOR);
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t full_bar[KB];
  __shared__ uint64_t done_bar;
  __shared__ uint32_t tmem_slot;
  __shared__ bool is_last;

  const int t = threadIdx.x, warp = t >> 5;
  const uint32_t smem = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int tiles_n = (p.N + BN - 1) / BN;
  const int tile = blockIdx.x, tile_m = tile / tiles_n, tile_n = tile - tile_m * tiles_n;
  const int z = blockIdx.z, splits = gridDim.y, split = blockIdx.y;
  const int m0 = tile_m * kT32BM, n0 = tile_n * BN;
  const int kblocks = (p.K + kT32BK - 1) / kT32BK;
  const int per = (kblocks + splits - 1) / splits;      // host guarantees per <= KB
  const int kb_beg = split * per, kb_end = min(kblocks, kb_beg + per);
  const int nkb = max(kb_end - kb_beg, 0);

  if (t == 0) {
    for (int s = 0; s < KB; ++s) mbar_init(&full_bar[s], 128);
    mbar_init(&done_bar, 1);
    fence_mbar_init();
  }
  if (warp == 4) t32_alloc<COLS>(&tmem_slot);
  t32_fence_before();
  __syncthreads();
  t32_fence_after();
  const uint32_t tmem = tmem_slot;

  if (warp < 4) {
    // ---- producers: everything in flight at once, one commit group per k-block
#pragma unroll 1
    for (int i = 0; i < nkb; ++i) {
      const int k0 = (kb_beg + i) * kT32BK;
      t32_stage<kT32BM, P::A_MN_MAJOR>(smem + i * A_TILE, m0, k0,
                                       [&](int r, int k, uint32_t dst) { p.a_chunk(z, r, k, dst); });
      t32_stage<BN, P::B_MN_MAJOR>(smem + KB * A_TILE + i * B_TILE, n0, k0,
                                   [&](int r, int k, uint32_t dst) { p.b_chunk(z, r, k, dst); });
      cp_commit();
    }
#pragma unroll 1
    for (int i = nkb; i < KB; ++i) cp_commit();     // empty groups keep the wait counts static
    t32_retire_groups<KB>(nkb, full_bar);
  } else if (t == 128) {
    // ---- MMA issuer
    const uint32_t idesc = t32_idesc(BN, P::A_MN_MAJOR, P::B_MN_MAJOR);
    for (int i = 0; i < nkb; ++i) {
      mbar_wait(&full_bar[i], 0);
      t32_fence_after();
      const uint32_t a_addr = smem + i * A_TILE, b_addr = smem + KB * A_TILE + i * B_TILE;
#pragma unroll
      for (int ks = 0; ks < kT32BK / 8; ++ks) {
        uint64_t ad = P::A_MN_MAJOR ? t32_desc(a_addr + ks * 1024, 4096, 512, 1) : t32_desc(a_addr + ks * 32, 16, 1024);
        uint64_t bd = P::B_MN_MAJOR ? t32_desc(b_addr + ks * 1024, 4096, 512, 1) : t32_desc(b_addr + ks * 32, 16, 1024);
        t32_mma(tmem, ad, bd, idesc, i > 0 || ks > 0);
      }
    }
    t32_commit(&done_bar);
  }

  if (warp < 4) {
    // ---- epilogue: thread = one output row (TMEM lane); 16 columns at a time
    if (nkb > 0) mbar_wait(&done_bar, 0);
    t32_fence_after();
    const int m = m0 + t;
    const uint32_t lane_addr = tmem + ((uint32_t)(warp * 32) << 16);
    const bool direct = splits == 1;
    float* mine = nullptr;
    if (!direct && !P::DEFER_REDUCE)
      mine = sk.ws + ((size_t)(z * gridDim.x + tile) * splits + split) * (kT32BM * BN);
#pragma unroll 1
    for (int c = 0; c < BN; c += 16) {
      float v[16];
      if (nkb > 0) t32_ld16(lane_addr + c, v);
      else {
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = 0.f;
      }
      const int valid = min(16, p.N - (n0 + c));
      if (P::DEFER_REDUCE) {
        if (m < p.M && valid > 0) p.store_partial(z, split, m, n0 + c, v, valid);
      } else if (direct) {
        if (m < p.M && valid > 0) p.store_row(z, m, n0 + c, v, valid);
      } else {
        // column-major partial tile: consecutive threads (rows) -> consecutive addresses
#pragma unroll
        for (int i = 0; i < 16; ++i) mine[(size_t)(c + i) * kT32BM + t] = v[i];
      }
    }
  }
  if (!P::DEFER_REDUCE && splits > 1) {
    // ---- deterministic split-K fix-up: last CTA of the tile adds the partials in split order
    __threadfence();
    __syncthreads();
    if (t == 0) {
      unsigned* cnt = sk.counters + z * gridDim.x + tile;
      unsigned old = atomicAdd(cnt, 1u);
      is_last = (old == (unsigned)splits - 1);
      if (is_last) *cnt = 0;
    }
    __syncthreads();
    if (is_last && warp < 4) {
      __threadfence();
      const int m = m0 + t;
      const float* base = sk.ws + (size_t)(z * gridDim.x + tile) * splits * (kT32BM * BN);
#pragma unroll 1
      for (int c = 0; c < BN; c += 16) {
        float v[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = 0.f;
        for (int q = 0; q < splits; ++q)
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] += __ldcg(base + (size_t)q * (kT32BM * BN) + (size_t)(c + i) * kT32BM + t);
        const int valid = min(16, p.N - (n0 + c));
        if (m < p.M && valid > 0) p.store_row(z, m, n0 + c, v, valid);
      }
    }
  }
  t32_fence_before();
  __syncthreads();
  if (warp == 4) t32_dealloc<COLS>(tmem);
}

template <class P>
inline int t32_splits_for(const P& p) {   // fewest splits that fit the slab
  int kblocks = (p.K + kT32BK - 1) / kT32BK;
  return (kblocks + P::KB_MAX - 1) / P::KB_MAX;
}
template <class P>
inline cudaError_t launch_gemm_tf32(const P& p, int Z, int splits, SplitK sk, cudaStream_t st) {
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
  int kblocks = (p.K + kT32BK - 1) / kT32BK;
  if ((kblocks + splits - 1) / splits > P::KB_MAX) return cudaErrorInvalidValue;
  int tiles = ((p.M + kT32BM - 1) / kT32BM) * ((p.N + P::BN - 1) / P::BN);
  dim3 grid(tiles, splits, Z);
  gemm_tf32<P><<<grid, kT32Threads, t32_smem_bytes<P>(), st>>>(p, sk);
  return cudaGetLastError();
}
template <class P>
inline size_t gemm_tf32_tiles(const P& p, int Z) {
  return (size_t)((p.M + kT32BM - 1) / kT32BM) * ((p.N + P::BN - 1) / P::BN) * Z;
}
template <class P>
inline size_t gemm_tf32_ws_floats(const P& p, int Z, int splits) {
  if (splits <= 1 || P::DEFER_REDUCE) return 0;
  return gemm_tf32_tiles(p, Z) * splits * kT32BM * P::BN;
}

}  // namespace dqnb
