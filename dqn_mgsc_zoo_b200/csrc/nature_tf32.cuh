// Operand providers that put every layer op of the Nature CNN (networks.py:181-221) on the
// TF32 tensor-core slab kernel gemm_tf32<P>.  Same GEMM formulations as nature_problems.cuh
// (implicit im2col forward, dgrad by stride-parity class, wgrad with a ones row for the bias),
// but operands move as 16-byte chunks (4 fp32) straight from the NHWC activations / HWIO
// weights into swizzled shared memory with cp.async; uint8 frames go through registers
// (0..255 is exact in TF32) and the 1/255 of networks.py:193 is applied to the accumulator.
// Index decoding is split into a per-row and a per-k context (see T32Map in gemm_tf32.cuh) so
// the divisions happen a handful of times per thread, not once per chunk.
// Weight gradients leave their split-K partials in gpart[split][...] (arena layout, bias row
// right behind the weights); rmsprop_reduce_kernel sums them in split order.
#pragma once
#include <cuda_fp16.h>

#include "gemm_tf32.cuh"
#include "head_kernels.cuh"
#include "nature_problems.cuh"

namespace dqnb {

// 4 packed uint8 -> 4 floats without I2F (a quarter-rate conversion: 32 K of them per conv1 tile were
// the producers' bottleneck): PRMT drops byte k into the mantissa of 2^23, one FADD removes the 2^23.
__device__ __forceinline__ float t32_u8_to_float(uint32_t w, uint32_t selector) {
  return __uint_as_float(__byte_perm(w, 0x4B000000u, selector)) - 8388608.0f;
}
__device__ __forceinline__ void t32_widen4(uint32_t dst, uint32_t w) {
  st16(dst, t32_u8_to_float(w, 0x7650), t32_u8_to_float(w, 0x7651), t32_u8_to_float(w, 0x7652),
       t32_u8_to_float(w, 0x7653));
}
// 4 floats -> 4 halves (round to nearest), one 8-byte store
__device__ __forceinline__ void t32_store_half4(__half* dst, const float4& o) {
  const __half2 lo = __floats2half2_rn(o.x, o.y), hi = __floats2half2_rn(o.z, o.w);
  uint2 u;
  u.x = *reinterpret_cast<const uint32_t*>(&lo);
  u.y = *reinterpret_cast<const uint32_t*>(&hi);
  *reinterpret_cast<uint2*>(dst) = u;
}
template <class T> struct InScale { static constexpr float value = 1.0f; };
template <> struct InScale<uint8_t> { static constexpr float value = 1.0f / 255.0f; };

// Contexts shared by most operands: a base pointer (null = out of range) on one side and an
// element offset (-1 = out of range) on the other.
template <class T>
__device__ __forceinline__ void t32_put_ptr_off(uint32_t dst, const T* base, int off, const T* fallback) {
  const bool ok = base != nullptr && off >= 0;
  cp16(dst, ok ? base + off : fallback, ok);
}
constexpr int kT32Ones = -2;   // offset code of the all-ones wgrad row (bias gradient)

// B operand = a row-major [K][N] fp32 matrix (haiku weights, or dy for the weight gradients),
// i.e. N-contiguous: MN-major chunks of 4 consecutive n for one k.
struct T32MatKN {
  using Row = int;            // n (first of 4) or -1
  using Col = const float*;   // mat + k*N or null
  __device__ static Row row(int n, int N) { return n < N ? n : -1; }
  __device__ static Col col(const float* mat, int k, int K, int N) { return k < K ? mat + (size_t)k * N : nullptr; }
};

// ---------------------------------------------------------------- conv forward
// C[m=(b,i,j), n=cout] = relu(scale * sum_k x[b, i*s+kh, j*s+kw, c] * W[k, n] + bias[n])
template <class L, class TIn>
struct T32ConvFwd {
  static constexpr const char* NAME = "conv_fwd";
  // conv1 (uint8 frames, 300 tiles for double-Q): half slabs so that two CTAs share an SM.
  static constexpr int BN = L::COUT, KB_MAX = sizeof(TIn) == 1 ? 4 : 8;
  static constexpr bool A_MN_MAJOR = false, B_MN_MAJOR = true, DEFER_REDUCE = false;
  static constexpr bool A_REG = sizeof(TIn) == 1;   // uint8 frames (conv1): 4 packed bytes per chunk
  static constexpr bool A_ROW16 = sizeof(TIn) == 1 && L::ROWLEN == 32;   // k-block = one 32-byte patch row
  static constexpr int B_RESIDENT_KB = (L::KCOL + kT32BK - 1) / kT32BK;   // stream kernel: the filters stay in smem
  int M, N, K;
  const TIn* x[kMaxNets];
  const float* w[kMaxNets];
  const float* bias[kMaxNets];
  float* y[kMaxNets];
  __half* yh[kMaxNets] = {};   // optional fp16 copy of the output (input of the next layer's fp16-operand forward)
  __device__ const uint8_t* a_row16(int z, int m) const { return reinterpret_cast<const uint8_t*>(a_row(z, m)); }
  __device__ int a_kb_off(int kb) const { return kb * (L::WIN * L::CIN); }
  using ARow = const TIn*;    // &x[b, i*s, j*s, 0] or null
  using ACol = int;           // offset of (kh, kw, c) inside the patch or -1
  using BRow = T32MatKN::Row;
  using BCol = T32MatKN::Col;
  __device__ ARow a_row(int z, int m) const {
    if (m >= M) return nullptr;
    int b = m / (L::HOUT * L::WOUT), r = m - b * (L::HOUT * L::WOUT);
    int i = r / L::WOUT, j = r - i * L::WOUT;
    return x[z] + ((size_t)(b * L::HIN + i * L::STRIDE) * L::WIN + j * L::STRIDE) * L::CIN;
  }
  __device__ ACol a_col(int, int k) const {
    if (k >= K) return -1;
    int kh = k / L::ROWLEN;
    return kh * (L::WIN * L::CIN) + (k - kh * L::ROWLEN);
  }
  __device__ void a_put(ARow r, ACol c, uint32_t dst) const { t32_put_ptr_off(dst, r, c, x[0]); }
  __device__ uint32_t a_get(ARow r, ACol c) const {
    return (r != nullptr && c >= 0) ? __ldg(reinterpret_cast<const uint32_t*>(reinterpret_cast<const uint8_t*>(r) + c)) : 0u;
  }
  __device__ void a_store(uint32_t dst, uint32_t wd) const { t32_widen4(dst, wd); }
  __device__ BRow b_row(int, int n) const { return T32MatKN::row(n, N); }
  __device__ BCol b_col(int z, int k) const { return T32MatKN::col(w[z], k, K, N); }
  __device__ void b_put(BRow r, BCol c, uint32_t dst) const { t32_put_ptr_off(dst, c, r, w[0]); }
  // slab kernel: destinations and bias are resolved while the MMAs run (gemm_tf32.cuh PRELOAD)
  static constexpr bool PRELOAD = true;
  struct Ctx { float* y; __half* yh; float4 b; };   // column-offset destinations (null: not wanted) and the bias
  struct Pre { int off; };                          // m * COUT
  __device__ Ctx store_ctx(int z, int, int n) const {
    return {y[z] != nullptr ? y[z] + n : nullptr, yh[z] != nullptr ? yh[z] + n : nullptr,
            __ldg(reinterpret_cast<const float4*>(bias[z] + n))};
  }
  __device__ Pre preload(const Ctx&, int, int, int m, int) const { return {m * L::COUT}; }
  __device__ void store4p(const Ctx& c, const Pre& q, float4 v) const {
    constexpr float s = InScale<TIn>::value;
    const float4 o = make_float4(fmaxf(fmaf(v.x, s, c.b.x), 0.f), fmaxf(fmaf(v.y, s, c.b.y), 0.f),
                                 fmaxf(fmaf(v.z, s, c.b.z), 0.f), fmaxf(fmaf(v.w, s, c.b.w), 0.f));
    if (c.y != nullptr) *reinterpret_cast<float4*>(c.y + q.off) = tf32_rn4(o);   // null: only the fp16 copy is consumed
    if (c.yh != nullptr) t32_store_half4(c.yh + q.off, o);
  }
  // (stream kernel: one call per float4 of a thread's row - kept free of the context bookkeeping above)
  __device__ void store4(int z, int, int m, int n, float4 v) const {
    const float4 b = __ldg(reinterpret_cast<const float4*>(bias[z] + n));
    constexpr float s = InScale<TIn>::value;
    const float4 o = make_float4(fmaxf(fmaf(v.x, s, b.x), 0.f), fmaxf(fmaf(v.y, s, b.y), 0.f),
                                 fmaxf(fmaf(v.z, s, b.z), 0.f), fmaxf(fmaf(v.w, s, b.w), 0.f));
    if (y[z] != nullptr) *reinterpret_cast<float4*>(y[z] + (size_t)m * N + n) = tf32_rn4(o);
    if (yh[z] != nullptr) t32_store_half4(yh[z] + (size_t)m * N + n, o);
  }
};

// Double-Q conv1: three nets but only two distinct inputs and two distinct weight sets.  Staging
// the im2col A tile is what a conv1 CTA spends its time on, so one GEMM covers all three: rows
// [0, Mh) are s_tm1, rows [Mh, 2Mh) are s_t, columns [0, COUT) the online filters and
// [COUT, 2 COUT) the target filters.  The (s_tm1, target) quadrant is computed and dropped (MMA time
// is not the limit); 200 tiles instead of 300.
template <class L, class TIn>
struct T32ConvFwdDoubleQ {
  static constexpr const char* NAME = "conv_fwd_dq";
  static constexpr int BN = 2 * L::COUT, KB_MAX = 8;
  static constexpr bool A_MN_MAJOR = false, B_MN_MAJOR = true, DEFER_REDUCE = false;
  static constexpr bool A_REG = sizeof(TIn) == 1;
  static constexpr bool A_ROW16 = sizeof(TIn) == 1 && L::ROWLEN == 32;
  static constexpr int B_RESIDENT_KB = (L::KCOL + kT32BK - 1) / kT32BK;
  __device__ const uint8_t* a_row16(int z, int m) const { return reinterpret_cast<const uint8_t*>(a_row(z, m)); }
  __device__ int a_kb_off(int kb) const { return kb * (L::WIN * L::CIN); }
  int M, N, K, Mh;          // M = 2 * Mh, N = 2 * COUT
  // (the two-entry arrays below are only ever read through selects: a dynamically indexed kernel
  // parameter array would be copied to local memory)
  const TIn* x[2];          // s_tm1, s_t
  const float* w[2];        // online, target
  const float* bias[2];
  float* y[kMaxNets];       // online(s_tm1), target(s_t), online(s_t)
  __half* yh[kMaxNets] = {};   // optional fp16 copies (see T32ConvFwd)
  using ARow = const TIn*;
  using ACol = int;
  struct BRow { int n, which; };   // n (first of 4) inside the net's COUT columns; n = -1: out of range
  using BCol = int;                // k or -1
  __device__ ARow a_row(int, int m) const {
    if (m >= M) return nullptr;
    const int half = m >= Mh, mm = m - half * Mh;
    int b = mm / (L::HOUT * L::WOUT), r = mm - b * (L::HOUT * L::WOUT);
    int i = r / L::WOUT, j = r - i * L::WOUT;
    return (half ? x[1] : x[0]) + ((size_t)(b * L::HIN + i * L::STRIDE) * L::WIN + j * L::STRIDE) * L::CIN;
  }
  __device__ ACol a_col(int, int k) const {
    if (k >= K) return -1;
    int kh = k / L::ROWLEN;
    return kh * (L::WIN * L::CIN) + (k - kh * L::ROWLEN);
  }
  __device__ void a_put(ARow r, ACol c, uint32_t dst) const { t32_put_ptr_off(dst, r, c, x[0]); }
  __device__ uint32_t a_get(ARow r, ACol c) const {
    return (r != nullptr && c >= 0) ? __ldg(reinterpret_cast<const uint32_t*>(reinterpret_cast<const uint8_t*>(r) + c)) : 0u;
  }
  __device__ void a_store(uint32_t dst, uint32_t wd) const { t32_widen4(dst, wd); }
  __device__ BRow b_row(int, int n) const {
    if (n >= N) return {-1, 0};
    return {n % L::COUT, n / L::COUT};
  }
  __device__ BCol b_col(int, int k) const { return k < K ? k : -1; }
  __device__ void b_put(const BRow& r, BCol c, uint32_t dst) const {
    const bool ok = r.n >= 0 && c >= 0;
    cp16(dst, ok ? (r.which ? w[1] : w[0]) + (size_t)c * L::COUT + r.n : w[0], ok);
  }
  static constexpr bool PRELOAD = true;
  // lo: rows < Mh (s_tm1), hi: rows >= Mh (s_t); the (s_tm1, target) quadrant has no destination at all
  struct Ctx { float *y_lo, *y_hi; __half *yh_lo, *yh_hi; float4 b; };
  struct Pre { int off; };    // (m mod Mh) * COUT, bit 31 set for the hi half
  __device__ Ctx store_ctx(int, int, int n) const {
    const int which = n >= L::COUT, nn = n - which * L::COUT;
    float* lo = which ? nullptr : y[0];
    float* hi = which ? y[1] : y[2];
    __half* lo_h = which ? nullptr : yh[0];
    __half* hi_h = which ? yh[1] : yh[2];
    return {lo != nullptr ? lo + nn : nullptr, hi != nullptr ? hi + nn : nullptr, lo_h != nullptr ? lo_h + nn : nullptr,
            hi_h != nullptr ? hi_h + nn : nullptr, __ldg(reinterpret_cast<const float4*>((which ? bias[1] : bias[0]) + nn))};
  }
  __device__ Pre preload(const Ctx&, int, int, int m, int) const {
    const int half = m >= Mh;
    return {((m - half * Mh) * L::COUT) | (half ? (int)0x80000000 : 0)};
  }
  __device__ void store4p(const Ctx& c, const Pre& q, float4 v) const {
    const bool half = q.off < 0;
    const int at = q.off & 0x7fffffff;
    float* dst = half ? c.y_hi : c.y_lo;
    __half* dsth = half ? c.yh_hi : c.yh_lo;
    constexpr float s = InScale<TIn>::value;
    const float4 o = make_float4(fmaxf(fmaf(v.x, s, c.b.x), 0.f), fmaxf(fmaf(v.y, s, c.b.y), 0.f),
                                 fmaxf(fmaf(v.z, s, c.b.z), 0.f), fmaxf(fmaf(v.w, s, c.b.w), 0.f));
    if (dst != nullptr) *reinterpret_cast<float4*>(dst + at) = tf32_rn4(o);   // null: only the fp16 copy is consumed
    if (dsth != nullptr) t32_store_half4(dsth + at, o);
  }
  // (stream kernel: one call per float4 of a thread's row - kept free of the context bookkeeping above)
  __device__ void store4(int, int, int m, int n, float4 v) const {
    const int half = m >= Mh, which = n >= L::COUT;
    if (!half && which) return;
    const int nn = n - which * L::COUT;
    float* dst = half ? (which ? y[1] : y[2]) : y[0];
    __half* dsth = half ? (which ? yh[1] : yh[2]) : yh[0];
    const float4 b = __ldg(reinterpret_cast<const float4*>((which ? bias[1] : bias[0]) + nn));
    constexpr float s = InScale<TIn>::value;
    const float4 o = make_float4(fmaxf(fmaf(v.x, s, b.x), 0.f), fmaxf(fmaf(v.y, s, b.y), 0.f),
                                 fmaxf(fmaf(v.z, s, b.z), 0.f), fmaxf(fmaf(v.w, s, b.w), 0.f));
    const size_t at = (size_t)(m - half * Mh) * L::COUT + nn;
    if (dst != nullptr) *reinterpret_cast<float4*>(dst + at) = tf32_rn4(o);
    if (dsth != nullptr) t32_store_half4(dsth + at, o);
  }
};

// The same GEMM with fp16 operands (slab kernel, small batches).  uint8 frames are exact in fp16 and
// fp16 has tf32's 11-bit significand, so nothing is lost against the tf32 path; the operand slab
// halves (96 KB: two CTAs per SM, so the 200 tiles of a batch-32 double-Q update run as ONE wave
// instead of two) and every MMA covers K = 16.  B is the learner's fp16 copy of both nets' conv1
// filters, transposed to K-major: w16[n][k], n < COUT online, n >= COUT target (conv1_h16_kernel).
// K counts 32-bit words here (KCOL / 2).
template <class L, class TIn>
struct H16ConvFwdDoubleQ : T32ConvFwdDoubleQ<L, TIn> {
  static constexpr int KB_MAX = 4;
  static constexpr bool A_MN_MAJOR = false, B_MN_MAJOR = false, DEFER_REDUCE = false;
  static constexpr bool F16 = true, A_ROW16H = true, A_ROW16 = false, A_REG = true;
  static constexpr int B_RESIDENT_KB = L::KCOL / 64;   // stream kernel: k-blocks of 64 halves
  const __half* w16;        // [2 * COUT][KCOL]
  static constexpr bool B_TMA = true;   // gemm_tf32.cuh: the filters by tensor-map TMA when a map is given
  const void* bmap = nullptr;           // device address of the CUtensorMap of w16, or null
  __device__ const void* b_tmap(int) const { return bmap; }
  using BRow = int;         // n or -1
  using BCol = int;         // word index k (2 halves each; first of 4) or -1
  __device__ BRow b_row(int, int n) const { return n < this->N ? n : -1; }
  __device__ BCol b_col(int, int k) const { return k < this->K ? k : -1; }
  __device__ void b_put(BRow r, BCol c, uint32_t dst) const {
    const bool ok = r >= 0 && c >= 0;
    cp16(dst, ok ? w16 + (size_t)r * L::KCOL + 2 * c : w16, ok);
  }
};
// conv2 / conv3 forward with fp16 operands (slab kernel, small batches): A = the previous layer's fp16 activation
// copy (NHWC, written by its epilogue next to the fp32 one the backward pass needs), B = the learner's fp16
// K-major filter copy w16[n][k].  Activations are post-ReLU values of O(1) and fp16 carries tf32's 11-bit
// significand, so the tolerance class is unchanged.  Half the operand bytes and half the MMAs: the whole K range
// (512 / 576 halves = 8 / 9 k-blocks) fits ONE slab, so the tiles run unsplit - no cluster, no DSMEM reduce.
// K counts 32-bit words (KCOL / 2).
template <class L>
struct H16ConvFwd {
  static constexpr const char* NAME = "conv_fwd_h16";
  static constexpr int BN = L::COUT, KB_MAX = (L::KCOL / 2 + kT32BK - 1) / kT32BK;
  static constexpr bool A_MN_MAJOR = false, B_MN_MAJOR = false, DEFER_REDUCE = false;
  static constexpr bool F16 = true;
  static constexpr bool EARLY_TRIGGER = true;   // conv2 -> conv3 -> fc1: 63 / 39 CTAs leave most SMs to the successor
  static constexpr int HOUT_WOUT = L::HOUT * L::WOUT, KWORDS = L::KCOL / 2;
  static_assert(L::ROWLEN % 8 == 0 && L::CIN % 8 == 0, "16-byte chunks of 8 halves must not straddle patch rows");
  int M, N, K;              // K = KCOL / 2 words
  const __half* x[kMaxNets];
  const __half* w16[kMaxNets];   // [COUT][KCOL]
  const float* bias[kMaxNets];
  float* y[kMaxNets];
  __half* yh[kMaxNets];     // may be null
  static constexpr bool B_TMA = true;   // gemm_tf32.cuh: the filters by tensor-map TMA when a map is given
  const void* bmap[kMaxNets] = {};      // device addresses of the CUtensorMaps of w16[z], or null
  __device__ const void* b_tmap(int z) const { return z == 0 ? bmap[0] : z == 1 ? bmap[1] : bmap[2]; }
  using ARow = const __half*;   // &x[b, i*s, j*s, 0] or null
  using ACol = int;             // offset (halves) of the chunk inside the patch or -1
  using BRow = const __half*;   // w16[z] + n * KCOL or null
  using BCol = int;             // offset (halves) inside the filter row or -1
  __device__ ARow a_row(int z, int m) const {
    if (m >= M) return nullptr;
    int b = m / (L::HOUT * L::WOUT), r = m - b * (L::HOUT * L::WOUT);
    int i = r / L::WOUT, j = r - i * L::WOUT;
    return x[z] + ((size_t)(b * L::HIN + i * L::STRIDE) * L::WIN + j * L::STRIDE) * L::CIN;
  }
  __device__ ACol a_col(int, int k) const {
    if (k >= K) return -1;
    const int kk = 2 * k, kh = kk / L::ROWLEN;
    return kh * (L::WIN * L::CIN) + (kk - kh * L::ROWLEN);
  }
  __device__ void a_put(ARow r, ACol c, uint32_t dst) const {
    const bool ok = r != nullptr && c >= 0;
    cp16(dst, ok ? r + c : x[0], ok);
  }
  __device__ BRow b_row(int z, int n) const { return n < N ? w16[z] + (size_t)n * L::KCOL : nullptr; }
  __device__ BCol b_col(int, int k) const { return k < K ? 2 * k : -1; }
  __device__ void b_put(BRow r, BCol c, uint32_t dst) const {
    const bool ok = r != nullptr && c >= 0;
    cp16(dst, ok ? r + c : w16[0], ok);
  }
  static constexpr bool PRELOAD = true;
  struct Ctx { float* y; __half* yh; float4 b; };
  struct Pre { int off; };
  __device__ Ctx store_ctx(int z, int, int n) const {
    return {y[z] != nullptr ? y[z] + n : nullptr, yh[z] != nullptr ? yh[z] + n : nullptr,
            __ldg(reinterpret_cast<const float4*>(bias[z] + n))};
  }
  __device__ Pre preload(const Ctx&, int, int, int m, int) const { return {m * L::COUT}; }
  __device__ void store4p(const Ctx& c, const Pre& q, float4 v) const {
    const float4 o = make_float4(fmaxf(v.x + c.b.x, 0.f), fmaxf(v.y + c.b.y, 0.f), fmaxf(v.z + c.b.z, 0.f),
                                 fmaxf(v.w + c.b.w, 0.f));
    if (c.y != nullptr) *reinterpret_cast<float4*>(c.y + q.off) = tf32_rn4(o);   // null: only the fp16 copy is consumed
    if (c.yh != nullptr) t32_store_half4(c.yh + q.off, o);
  }
  __device__ void store4(int z, int split, int m, int n, float4 v) const {
    const Ctx c = store_ctx(z, split, n);
    store4p(c, preload(c, z, split, m, n), v);
  }
};
// w[k][n] fp32 (one net's conv1 filters, HWIO) -> dst[n][k] fp16
__global__ void conv1_h16_kernel(const float* __restrict__ w, __half* __restrict__ dst, int K, int N) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < K * N) dst[(size_t)(i % N) * K + i / N] = __float2half_rn(w[i]);
}

// src fp32 -> dst fp16, same layout (n a multiple of 4)
__global__ void half_copy_kernel(const float* __restrict__ src, __half* __restrict__ dst, int64_t n) {
  const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i < n) t32_store_half4(dst + i, __ldg(reinterpret_cast<const float4*>(src + i)));
}

// ---------------------------------------------------------------- conv weight gradient
// dW[m'=(kh,kw,c), n] = scale * sum_{k'=(b,i,j)} x[...] * dy[k', n]; row m' == KCOL is all ones
// and yields the bias gradient, which sits right behind the weights in the arena.  Where KCOL is
// a multiple of 128 that one row would cost a whole extra M tile (conv1: 3 tiles instead of 2),
// so BIAS_ROW = false leaves it to bias_colsum_kernel, which fills the same gpart row.
template <class L, class TIn, int KB_, bool BIAS_ROW = true>
struct T32ConvWgrad {
  static constexpr const char* NAME = "conv_wgrad";
  static constexpr int BN = L::COUT, KB_MAX = KB_;
  static constexpr bool A_MN_MAJOR = true, B_MN_MAJOR = true, DEFER_REDUCE = true;
  static constexpr bool A_REG = sizeof(TIn) == 1;
  static constexpr bool B_REG = sizeof(TIn) == 1 && L::COUT == 32;   // conv1 (stream kernel): dy rides the register prefetch too
  // conv1 (stream kernel): both operands ride the in-order cp.async byte ring (gemm_tf32_stream.cuh)
  static constexpr bool A_COL16 = sizeof(TIn) == 1 && L::ROWLEN == 32 && L::COUT == 32 && !BIAS_ROW;
  __device__ const uint8_t* a_src16(int m0, int k, int c16) const {
    ACol c = a_col(0, k);
    if (c == nullptr) return nullptr;
    return reinterpret_cast<const uint8_t*>(c) + (m0 / L::ROWLEN + (c16 >> 1)) * (L::WIN * L::CIN) + (c16 & 1) * 16;
  }
  __device__ const float* b_src(int r, const float* c) const { return (c != nullptr && r >= 0) ? c + r : nullptr; }
  static constexpr int ROWS = L::KCOL + (BIAS_ROW ? 1 : 0);
  static constexpr size_t PART = (size_t)(L::KCOL + 1) * L::COUT;   // floats per split in gpart
  int M, N, K;              // M = ROWS, N = COUT, K = B*HOUT*WOUT
  const TIn* x;
  const float* dy;
  float* gpart;             // [splits][(KCOL + 1) * COUT]
  using ARow = int;           // offset of (kh, kw, c..c+3) inside the patch, kT32Ones, or -1
  using ACol = const TIn*;    // &x[b, i*s, j*s, 0] of output position k, or null
  using BRow = T32MatKN::Row;
  using BCol = T32MatKN::Col;
  __device__ ARow a_row(int, int m) const {
    if (m >= ROWS) return -1;
    if (m == L::KCOL) return kT32Ones;
    int kh = m / L::ROWLEN;
    return kh * (L::WIN * L::CIN) + (m - kh * L::ROWLEN);
  }
  __device__ ACol a_col(int, int k) const {
    if (k >= K) return nullptr;
    int b = k / (L::HOUT * L::WOUT), q = k - b * (L::HOUT * L::WOUT);
    int i = q / L::WOUT, j = q - i * L::WOUT;
    return x + ((size_t)(b * L::HIN + i * L::STRIDE) * L::WIN + j * L::STRIDE) * L::CIN;
  }
  __device__ void a_put(ARow r, ACol c, uint32_t dst) const {
    if (r == kT32Ones) st16(dst, c != nullptr ? 1.f : 0.f, 0.f, 0.f, 0.f);
    else t32_put_ptr_off(dst, c, r, x);
  }
  __device__ uint32_t a_get(ARow r, ACol c) const {
    if (c == nullptr || r == -1) return 0u;
    if (r == kT32Ones) return 1u;       // widens to (1, 0, 0, 0)
    return __ldg(reinterpret_cast<const uint32_t*>(reinterpret_cast<const uint8_t*>(c) + r));
  }
  __device__ void a_store(uint32_t dst, uint32_t wd) const { t32_widen4(dst, wd); }
  __device__ BRow b_row(int, int n) const { return T32MatKN::row(n, N); }
  __device__ BCol b_col(int, int k) const { return T32MatKN::col(dy, k, K, N); }
  __device__ void b_put(BRow r, BCol c, uint32_t dst) const { t32_put_ptr_off(dst, c, r, dy); }
  __device__ float4 b_get(BRow r, BCol c) const {
    return (c != nullptr && r >= 0) ? __ldg(reinterpret_cast<const float4*>(c + r)) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  __device__ void store4(int, int split, int m, int n, float4 v) const {
    const float s = m < L::KCOL ? InScale<TIn>::value : 1.0f;
    *reinterpret_cast<float4*>(gpart + (size_t)split * PART + (size_t)m * N + n) =
        make_float4(v.x * s, v.y * s, v.z * s, v.w * s);
  }
};

// Bias gradient of a conv layer = column sums of dy[rows][C]: block q adds its share of the rows
// (fixed order) into gpart[q][KCOL*C + c], next to split q's weight-gradient partial.
template <int C>
__global__ void __launch_bounds__(256) bias_colsum_kernel(const float* __restrict__ dy, int rows, float* __restrict__ gpart,
                                                          size_t part_stride, size_t row_off, unsigned long long* tl) {
  const unsigned long long tl_t0 = tl_begin(tl);
  constexpr int LANES = 256 / C;
  __shared__ float acc_s[LANES][C];
  const int q = blockIdx.x, nq = gridDim.x, t = threadIdx.x, c = t % C, lane = t / C;
  const int beg = (int)((int64_t)rows * q / nq), end = (int)((int64_t)rows * (q + 1) / nq);
  float acc = 0.f;
#pragma unroll 8
  for (int r = beg + lane; r < end; r += LANES) acc += __ldg(dy + (size_t)r * C + c);
  acc_s[lane][c] = acc;
  __syncthreads();
  if (t < C) {
    float v = 0.f;
#pragma unroll
    for (int l = 0; l < LANES; ++l) v += acc_s[l][t];
    gpart[(size_t)q * part_stride + row_off + t] = v;
  }
  tl_end(tl, tl_t0);
}

// An empty node: delays the side-branch kernel queued behind it by one launch latency, so that the
// chain kernel released by the same event gets its CTAs placed first (launch priorities order
// pending CTAs but cannot evict ones that are already resident).
__global__ void stagger_kernel() {}

#include "nature_tf32_more.inl"

}  // namespace dqnb
