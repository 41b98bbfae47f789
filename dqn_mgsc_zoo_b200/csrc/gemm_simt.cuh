// Generic fp32 SIMT tile GEMM with functor operand loaders (implicit im2col etc.) and a
// deterministic split-K fix-up.  This is the fp32 / rel-1e-5 precision mode of the learner
// (K2/K3 "SIMT twin"); the tensor-core mode lives in gemm_tf32.cuh / gemm_tf32_stream.cuh.
//
// A problem type P provides:
//   static constexpr int BM, BN;            tile (multiples of 16; 256 threads = 16x16)
//   static constexpr bool A_K_CONTIG, B_N_CONTIG;   which index is contiguous in memory
//   int M, N, K;                            GEMM extents (C[M,N] = sum_k A[m,k] B[k,n])
//   typename ARow / BCol;                   per-row / per-column precomputed context
//   ARow a_row(int z, int m);  float a(const ARow&, int k);      (must bounds-check m, k)
//   BCol b_col(int z, int n);  float b(const BCol&, int k);      (must bounds-check n, k)
//   void store(int z, int m, int n, float acc);                  (only called in range)
// grid = (tiles_m * tiles_n, splits, Z).  With splits > 1 each CTA writes its partial tile to
// `ws`, and the last CTA to arrive on the tile's counter adds the partials in split order
// (bit-reproducible) and calls store(); counters reset themselves for the next launch.
#pragma once
#include "common.cuh"

namespace dqnb {

constexpr int kGemmThreads = 256;
constexpr int kBK = 16;

struct SplitK {
  float* ws;          // [Z * tiles * splits][BM*BN]
  unsigned* counters; // [Z * tiles], zero before first use
  long long* dbg = nullptr;   // optional: CTA (0,0,0) writes clock64() at its phase boundaries (gemm_tf32)
  unsigned long long* tl = nullptr;   // optional timeline slot (common.cuh tl_begin / tl_end)
};

template <class P>
__global__ void __launch_bounds__(kGemmThreads) gemm_simt(P p, SplitK sk) {
  constexpr int BM = P::BM, BN = P::BN, TM = BM / 16, TN = BN / 16;
  constexpr int AE = BM * kBK / kGemmThreads, BE = BN * kBK / kGemmThreads;
  __shared__ __align__(16) float As[kBK][BM + 4];
  __shared__ __align__(16) float Bs[kBK][BN + 4];
  __shared__ bool is_last;
  tl_begin(sk.tl);

  const int t = threadIdx.x, ty = t >> 4, tx = t & 15;
  const int tiles_n = (p.N + BN - 1) / BN;
  const int tile = blockIdx.x, tile_m = tile / tiles_n, tile_n = tile - tile_m * tiles_n;
  const int z = blockIdx.z, splits = gridDim.y, split = blockIdx.y;
  const int m0 = tile_m * BM, n0 = tile_n * BN;
  // K range of this split, in multiples of kBK
  const int kchunks = (p.K + kBK - 1) / kBK;
  const int per = (kchunks + splits - 1) / splits;
  const int kbeg = split * per * kBK, kend = min(p.K, (split + 1) * per * kBK);

  // operand element -> thread mapping (contiguous index fastest across the warp)
  int a_m[AE], a_k[AE], b_n[BE], b_k[BE];
  typename P::ARow arow[AE];
  typename P::BCol bcol[BE];
#pragma unroll
  for (int i = 0; i < AE; ++i) {
    int e = t + i * kGemmThreads;
    if (P::A_K_CONTIG) { a_k[i] = e % kBK; a_m[i] = e / kBK; }
    else               { a_m[i] = e % BM;  a_k[i] = e / BM; }
    arow[i] = p.a_row(z, m0 + a_m[i]);
  }
#pragma unroll
  for (int i = 0; i < BE; ++i) {
    int e = t + i * kGemmThreads;
    if (P::B_N_CONTIG) { b_n[i] = e % BN;  b_k[i] = e / BN; }
    else               { b_k[i] = e % kBK; b_n[i] = e / kBK; }
    bcol[i] = p.b_col(z, n0 + b_n[i]);
  }

  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

  float ra[AE], rb[BE];
  if (kbeg < kend) {
#pragma unroll
    for (int i = 0; i < AE; ++i) ra[i] = p.a(arow[i], kbeg + a_k[i]);
#pragma unroll
    for (int i = 0; i < BE; ++i) rb[i] = p.b(bcol[i], kbeg + b_k[i]);
  }
  for (int k0 = kbeg; k0 < kend; k0 += kBK) {
#pragma unroll
    for (int i = 0; i < AE; ++i) As[a_k[i]][a_m[i]] = ra[i];
#pragma unroll
    for (int i = 0; i < BE; ++i) Bs[b_k[i]][b_n[i]] = rb[i];
    __syncthreads();
    if (k0 + kBK < kend) {   // prefetch next tile into registers while computing this one
#pragma unroll
      for (int i = 0; i < AE; ++i) ra[i] = p.a(arow[i], k0 + kBK + a_k[i]);
#pragma unroll
      for (int i = 0; i < BE; ++i) rb[i] = p.b(bcol[i], k0 + kBK + b_k[i]);
    }
#pragma unroll
    for (int k = 0; k < kBK; ++k) {
      float av[TM], bv[TN];
#pragma unroll
      for (int i = 0; i < TM; ++i) av[i] = As[k][ty * TM + i];
#pragma unroll
      for (int j = 0; j < TN; ++j) bv[j] = Bs[k][tx * TN + j];
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }

  if (splits == 1) {
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
      for (int j = 0; j < TN; ++j) {
        int m = m0 + ty * TM + i, n = n0 + tx * TN + j;
        if (m < p.M && n < p.N) p.store(z, m, n, acc[i][j]);
      }
    tl_end(sk.tl);
    return;
  }
  // ---- deterministic split-K: partial -> workspace; last arriver reduces in split order
  const int tiles = gridDim.x;
  float* mine = sk.ws + ((size_t)(z * tiles + tile) * splits + split) * (BM * BN);
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) mine[(ty * TM + i) * BN + tx * TN + j] = acc[i][j];
  __threadfence();
  __syncthreads();
  if (t == 0) {
    unsigned* c = sk.counters + z * tiles + tile;
    unsigned old = atomicAdd(c, 1u);
    is_last = (old == (unsigned)splits - 1);
    if (is_last) *c = 0;
  }
  __syncthreads();
  if (!is_last) {
    tl_end(sk.tl);
    return;
  }
  __threadfence();
  const float* base = sk.ws + (size_t)(z * tiles + tile) * splits * (BM * BN);
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      int idx = (ty * TM + i) * BN + tx * TN + j;
      float s = 0.f;
      for (int q = 0; q < splits; ++q) s += __ldcg(base + (size_t)q * (BM * BN) + idx);
      int m = m0 + ty * TM + i, n = n0 + tx * TN + j;
      if (m < p.M && n < p.N) p.store(z, m, n, s);
    }
  tl_end(sk.tl);
}

template <class P>
inline cudaError_t launch_gemm(const P& p, int Z, int splits, SplitK sk, cudaStream_t st) {
  int tiles = ((p.M + P::BM - 1) / P::BM) * ((p.N + P::BN - 1) / P::BN);
  dim3 grid(tiles, splits, Z);
  gemm_simt<P><<<grid, kGemmThreads, 0, st>>>(p, sk);
  return cudaGetLastError();
}

template <class P>
inline size_t gemm_ws_floats(const P& p, int Z, int splits) {
  if (splits <= 1) return 0;
  size_t tiles = (size_t)((p.M + P::BM - 1) / P::BM) * ((p.N + P::BN - 1) / P::BN);
  return tiles * Z * splits * P::BM * P::BN;
}
template <class P>
inline size_t gemm_tiles(const P& p, int Z) {
  return (size_t)((p.M + P::BM - 1) / P::BM) * ((p.N + P::BN - 1) / P::BN) * Z;
}

}  // namespace dqnb
