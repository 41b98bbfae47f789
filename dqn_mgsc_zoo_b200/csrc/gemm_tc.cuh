// tcgen05 (5th-gen tensor core) tile GEMM with thread-staged operands and TMEM accumulators.
//
// C[M,N] (+)= A[M,K] * B[N,K]^T, bf16 operands, fp32 accumulate.  One CTA = one 128 x BN output
// tile over a K range (split-K as in gemm_simt.cuh).  The operands are *gathered by the CTA's
// threads* (implicit im2col, strided dgrad taps, uint8->bf16, fp32->bf16 happen here) into the
// canonical 128-byte-swizzled shared-memory layouts tcgen05.mma consumes, 16 bytes (8 bf16) at
// a time; one elected thread issues the MMAs; tcgen05.commit on an mbarrier releases each smem
// stage; the epilogue reads the accumulator tile back with tcgen05.ld (one TMEM lane = one row).
//
// A problem type P provides
//   static constexpr int BN;                      UMMA N (16..256, multiple of 16)
//   static constexpr bool A_MN_MAJOR, B_MN_MAJOR; operand source contiguous along M/N instead of K
//   int M, N, K;
//   uint4 a_chunk(int z, int i, int j) const;     K-major: 8 bf16 A[m=i, k=j..j+7]
//                                                 MN-major: 8 bf16 A[m=i..i+7, k=j]   (zero-fill OOB)
//   uint4 b_chunk(int z, int i, int j) const;     same for B[n, k]
//   void store_row(int z, int m, int n0, const float (&v)[16], int valid_n) const;  16 columns of row m
// grid = (tiles_m * tiles_n, splits, Z), block = 128 threads.
#pragma once
#include <cuda_bf16.h>

#include "gemm_simt.cuh"

namespace dqnb {

constexpr int kTcThreads = 128;
constexpr int kTcBM = 128;
constexpr int kTcBK = 64;      // bf16 per k-block = one 128-byte swizzle row
constexpr int kTcStages = 3;

// ---- descriptors (bit layouts: cute/arch/mma_sm100_desc.hpp) ----
// smem matrix descriptor, SWIZZLE_128B, version 1.  K-major: LBO field 1, SBO = 1024 B between
// 8-row groups.  MN-major: LBO = bytes between 64-element MN blocks, SBO = 1024 B between 8-k groups.
__device__ __forceinline__ uint64_t tc_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;    // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;    // LayoutType::SWIZZLE_128B
  return d;
}
// instruction descriptor for kind::f16: D=f32, A=B=bf16, M=128.
__device__ __forceinline__ uint32_t tc_idesc(int n, bool a_mn, bool b_mn) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((a_mn ? 1u : 0u) << 15) | ((b_mn ? 1u : 0u) << 16) |
         ((uint32_t)(n >> 3) << 17) | ((uint32_t)(kTcBM >> 4) << 24);
}

__device__ __forceinline__ void tc_mma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, bool acc) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"((uint32_t)acc)
      : "memory");
}
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

template <int kCols>
__device__ __forceinline__ void tc_alloc(uint32_t* smem_slot) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)), "n"(kCols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int kCols>
__device__ __forceinline__ void tc_dealloc(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols) : "memory");
}
// 16 consecutive fp32 columns of this thread's TMEM lane.
__device__ __forceinline__ void tc_ld16(uint32_t taddr, float (&v)[16]) {
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

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ uint4 pack8_bf16(const float (&f)[8]) {
  return make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
}

__host__ __device__ constexpr int tc_tmem_cols(int n) { return n <= 32 ? 32 : n <= 64 ? 64 : n <= 128 ? 128 : 256; }
template <class P>
__host__ __device__ constexpr int tc_b_bytes() {   // smem bytes of one B stage
  return P::B_MN_MAJOR ? ((P::BN + 63) / 64) * (kTcBK / 8) * 1024 : P::BN * 128;
}
template <class P>
__host__ __device__ constexpr int tc_smem_bytes() { return kTcStages * (kTcBM * 128 + tc_b_bytes<P>()) + 1024; }

// Stage one [ROWS x 64] operand k-block into smem.  K-major: tile[row][k]; MN-major: tile[k][mn].
template <int ROWS, bool MN_MAJOR, class F>
__device__ __forceinline__ void tc_stage(uint8_t* tile, int r0, int k0, F&& chunk) {
  if (!MN_MAJOR) {
    // chunk id = row * 8 + c ; consecutive threads take consecutive k-chunks of a row
    for (int id = threadIdx.x; id < ROWS * 8; id += kTcThreads) {
      int row = id >> 3, c = id & 7;
      uint4 v = chunk(r0 + row, k0 + c * 8);
      *reinterpret_cast<uint4*>(tile + (row >> 3) * 1024 + (row & 7) * 128 + ((c ^ (row & 7)) << 4)) = v;
    }
  } else {
    constexpr int CH = ((ROWS + 63) / 64) * 8;       // 16-byte mn-chunks per k row
    constexpr int LBO = (kTcBK / 8) * 1024;
    for (int id = threadIdx.x; id < kTcBK * CH; id += kTcThreads) {
      int k = id / CH, cc = id - k * CH;
      uint4 v = cc * 8 < ROWS ? chunk(r0 + cc * 8, k0 + k) : make_uint4(0, 0, 0, 0);
      int j = cc >> 3, c = cc & 7, g = k >> 3, kr = k & 7;
      *reinterpret_cast<uint4*>(tile + j * LBO + g * 1024 + kr * 128 + ((c ^ kr) << 4)) = v;
    }
  }
}

template <class P>
__global__ void __launch_bounds__(kTcThreads) gemm_tc(P p, SplitK sk) {
  constexpr int BN = P::BN, COLS = tc_tmem_cols(BN);
  constexpr int A_BYTES = kTcBM * 128, B_BYTES = tc_b_bytes<P>();
  constexpr int MN_LBO = (kTcBK / 8) * 1024;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t empty_bar[kTcStages];
  __shared__ uint64_t done_bar;
  __shared__ uint32_t tmem_slot;
  __shared__ bool is_last;

  const int t = threadIdx.x, warp = t >> 5;
  const int tiles_n = (p.N + BN - 1) / BN;
  const int tile = blockIdx.x, tile_m = tile / tiles_n, tile_n = tile - tile_m * tiles_n;
  const int z = blockIdx.z, splits = gridDim.y, split = blockIdx.y;
  const int m0 = tile_m * kTcBM, n0 = tile_n * BN;
  const int kblocks = (p.K + kTcBK - 1) / kTcBK;
  const int per = (kblocks + splits - 1) / splits;
  const int kb_beg = split * per, kb_end = min(kblocks, (split + 1) * per);

  if (t == 0) {
    for (int s = 0; s < kTcStages; ++s) mbar_init(&empty_bar[s], 1);
    mbar_init(&done_bar, 1);
    fence_mbar_init();
  }
  if (warp == 0) tc_alloc<COLS>(&tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  const uint32_t idesc = tc_idesc(BN, P::A_MN_MAJOR, P::B_MN_MAJOR);

  for (int kb = kb_beg; kb < kb_end; ++kb) {
    const int it = kb - kb_beg, s = it % kTcStages;
    uint8_t* a_tile = smem + s * (A_BYTES + B_BYTES);
    uint8_t* b_tile = a_tile + A_BYTES;
    if (it >= kTcStages) mbar_wait(&empty_bar[s], ((it / kTcStages) - 1) & 1);   // MMAs done with this stage
    tc_stage<kTcBM, P::A_MN_MAJOR>(a_tile, m0, kb * kTcBK, [&](int i, int j) { return p.a_chunk(z, i, j); });
    tc_stage<BN, P::B_MN_MAJOR>(b_tile, n0, kb * kTcBK, [&](int i, int j) { return p.b_chunk(z, i, j); });
    fence_proxy_async_smem();      // generic-proxy smem writes -> visible to the tensor core (async proxy)
    __syncthreads();
    if (t == 0) {
      tc_fence_after();
      const uint32_t a_addr = smem_u32(a_tile), b_addr = smem_u32(b_tile);
#pragma unroll
      for (int ks = 0; ks < kTcBK / 16; ++ks) {
        uint64_t ad = P::A_MN_MAJOR ? tc_desc(a_addr + ks * 2048, MN_LBO, 1024) : tc_desc(a_addr + ks * 32, 16, 1024);
        uint64_t bd = P::B_MN_MAJOR ? tc_desc(b_addr + ks * 2048, MN_LBO, 1024) : tc_desc(b_addr + ks * 32, 16, 1024);
        tc_mma(tmem, ad, bd, idesc, it > 0 || ks > 0);
      }
      tc_commit(&empty_bar[s]);
    }
  }
  if (t == 0) tc_commit(&done_bar);
  if (kb_end > kb_beg) mbar_wait(&done_bar, 0);
  tc_fence_after();

  // ---- epilogue: thread = one output row (TMEM lane); 16 columns at a time
  const int m = m0 + t;
  const uint32_t lane_addr = tmem + ((uint32_t)(warp * 32) << 16);
  if (splits == 1) {
#pragma unroll 1
    for (int c = 0; c < BN; c += 16) {
      float v[16];
      if (kb_end > kb_beg) tc_ld16(lane_addr + c, v);
      else {
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = 0.f;
      }
      int valid = min(16, p.N - (n0 + c));
      if (m < p.M && valid > 0) p.store_row(z, m, n0 + c, v, valid);
    }
  } else {
    const int tiles = gridDim.x;
    float* mine = sk.ws + ((size_t)(z * tiles + tile) * splits + split) * (kTcBM * BN);
#pragma unroll 1
    for (int c = 0; c < BN; c += 16) {
      float v[16];
      if (kb_end > kb_beg) tc_ld16(lane_addr + c, v);
      else {
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = 0.f;
      }
      // column-major partial tile: consecutive threads (rows) -> consecutive addresses
#pragma unroll
      for (int i = 0; i < 16; ++i) mine[(size_t)(c + i) * kTcBM + t] = v[i];
    }
    __threadfence();
    __syncthreads();
    if (t == 0) {
      unsigned* cnt = sk.counters + z * tiles + tile;
      unsigned old = atomicAdd(cnt, 1u);
      is_last = (old == (unsigned)splits - 1);
      if (is_last) *cnt = 0;
    }
    __syncthreads();
    if (is_last) {
      __threadfence();
      const float* base = sk.ws + (size_t)(z * tiles + tile) * splits * (kTcBM * BN);
#pragma unroll 1
      for (int c = 0; c < BN; c += 16) {
        float v[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = 0.f;
        for (int q = 0; q < splits; ++q)
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] += __ldcg(base + (size_t)q * (kTcBM * BN) + (size_t)(c + i) * kTcBM + t);
        int valid = min(16, p.N - (n0 + c));
        if (m < p.M && valid > 0) p.store_row(z, m, n0 + c, v, valid);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tc_dealloc<COLS>(tmem);
}

template <class P>
inline cudaError_t launch_gemm_tc(const P& p, int Z, int splits, SplitK sk, cudaStream_t st) {
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(gemm_tc<P>, cudaFuncAttributeMaxDynamicSharedMemorySize, tc_smem_bytes<P>());
    if (e != cudaSuccess) return e;
    configured = true;
  }
  int tiles = ((p.M + kTcBM - 1) / kTcBM) * ((p.N + P::BN - 1) / P::BN);
  dim3 grid(tiles, splits, Z);
  gemm_tc<P><<<grid, kTcThreads, tc_smem_bytes<P>(), st>>>(p, sk);
  return cudaGetLastError();
}
template <class P>
inline size_t gemm_tc_ws_floats(const P& p, int Z, int splits) {
  if (splits <= 1) return 0;
  size_t tiles = (size_t)((p.M + kTcBM - 1) / kTcBM) * ((p.N + P::BN - 1) / P::BN);
  return tiles * Z * splits * kTcBM * P::BN;
}
template <class P>
inline size_t gemm_tc_tiles(const P& p, int Z) {
  return (size_t)((p.M + kTcBM - 1) / kTcBM) * ((p.N + P::BN - 1) / P::BN) * Z;
}

}  // namespace dqnb
