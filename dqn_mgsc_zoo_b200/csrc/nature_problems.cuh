// Operand loaders that turn every layer op of the Nature CNN (networks.py:181-221) into a
// GEMM for gemm_simt<P>: implicit-im2col forward conv, conv wgrad (+bias row), conv dgrad by
// stride-parity class, and the dense (fc1) forward / wgrad / dgrad.  Activations are NHWC,
// weights HWIO = row-major [K = kh*kw*cin, N = cout], exactly haiku's layout, so the flat
// parameter arena needs no repacking.
#pragma once
#include "gemm_simt.cuh"

namespace dqnb {

constexpr int kMaxNets = 3;   // online(s_tm1), target(s_t), online(s_t) [double-Q]

template <int HIN_, int WIN_, int CIN_, int KS_, int STRIDE_, int COUT_>
struct ConvShape {
  static constexpr int HIN = HIN_, WIN = WIN_, CIN = CIN_, KS = KS_, STRIDE = STRIDE_, COUT = COUT_;
  static constexpr int HOUT = (HIN - KS) / STRIDE + 1, WOUT = (WIN - KS) / STRIDE + 1;
  static constexpr int KCOL = KS * KS * CIN;          // im2col row length
  static constexpr int ROWLEN = KS * CIN;             // contiguous run inside one kernel row
  static constexpr int IN_PER_IMG = HIN * WIN * CIN, OUT_PER_IMG = HOUT * WOUT * COUT;
};
using Conv1 = ConvShape<84, 84, 4, 8, 4, 32>;
using Conv2 = ConvShape<20, 20, 32, 4, 2, 64>;
using Conv3 = ConvShape<9, 9, 64, 3, 1, 64>;
constexpr int kFc1In = Conv3::OUT_PER_IMG;   // 3136
constexpr int kFc1Out = 512;

__device__ __forceinline__ float to_unit(uint8_t v) { return (float)v * (1.0f / 255.0f); }  // x/255 (networks.py:193)
__device__ __forceinline__ float to_unit(float v) { return v; }

// ---------------------------------------------------------------- conv forward
// C[m=(b,i,j), n=cout] = relu(sum_k x[b, i*s+kh, j*s+kw, c] * W[k, n] + bias[n])
template <class L, class TIn, int BM_>
struct ConvFwd {
  static constexpr const char* NAME = "conv_fwd";
  static constexpr int BM = BM_, BN = (L::COUT < 64 ? L::COUT : 64);
  static constexpr bool A_K_CONTIG = true, B_N_CONTIG = true;
  int M, N, K;
  const TIn* x[kMaxNets];
  const float* w[kMaxNets];
  const float* bias[kMaxNets];
  float* y[kMaxNets];
  struct ARow { const TIn* p; };
  struct BCol { const float* p; };
  __device__ ARow a_row(int z, int m) const {
    if (m >= M) return {nullptr};
    int b = m / (L::HOUT * L::WOUT), r = m - b * (L::HOUT * L::WOUT);
    int i = r / L::WOUT, j = r - i * L::WOUT;
    return {x[z] + ((size_t)(b * L::HIN + i * L::STRIDE) * L::WIN + j * L::STRIDE) * L::CIN};
  }
  __device__ float a(const ARow& r, int k) const {
    if (!r.p || k >= K) return 0.f;
    int kh = k / L::ROWLEN;
    return to_unit(r.p[kh * (L::WIN * L::CIN) + (k - kh * L::ROWLEN)]);
  }
  __device__ BCol b_col(int z, int n) const { return {n < N ? w[z] + n : nullptr}; }
  __device__ float b(const BCol& c, int k) const { return (c.p && k < K) ? c.p[(size_t)k * N] : 0.f; }
  __device__ void store(int z, int m, int n, float acc) const {
    const float v = acc + bias[z][n];
    y[z][(size_t)m * N + n] = linear ? v : fmaxf(v, 0.f);
  }
  int linear = 0;           // 1: no ReLU (the layer-local products of the MGSC adjoint dots)
};

// ---------------------------------------------------------------- conv weight gradient
// dW[m'=(kh,kw,c), n] = sum_{k'=(b,i,j)} x[...] * dy[k', n];  row m' == KCOL is all ones and
// yields the bias gradient, which sits right behind the weights in the arena.
template <class L, class TIn>
struct ConvWgrad {
  static constexpr const char* NAME = "conv_wgrad";
  static constexpr int BM = 64, BN = (L::COUT < 64 ? L::COUT : 64);
  static constexpr bool A_K_CONTIG = false, B_N_CONTIG = true;
  int M, N, K;              // M = KCOL + 1, N = COUT, K = B*HOUT*WOUT
  const TIn* x;
  const float* dy;
  float* grad;              // -> [KCOL*COUT weights][COUT bias]
  struct ARow { int off; };   // offset of (kh,kw,c) inside a patch; -1 ones row; -2 out of range
  struct BCol { const float* p; };
  __device__ ARow a_row(int, int m) const {
    if (m > L::KCOL) return {-2};
    if (m == L::KCOL) return {-1};
    int kh = m / L::ROWLEN;
    return {kh * (L::WIN * L::CIN) + (m - kh * L::ROWLEN)};
  }
  __device__ float a(const ARow& r, int k) const {
    if (r.off == -2 || k >= K) return 0.f;
    if (r.off == -1) return 1.f;
    int b = k / (L::HOUT * L::WOUT), q = k - b * (L::HOUT * L::WOUT);
    int i = q / L::WOUT, j = q - i * L::WOUT;
    return to_unit(x[((size_t)(b * L::HIN + i * L::STRIDE) * L::WIN + j * L::STRIDE) * L::CIN + r.off]);
  }
  __device__ BCol b_col(int, int n) const { return {n < N ? dy + n : nullptr}; }
  __device__ float b(const BCol& c, int k) const { return (c.p && k < K) ? c.p[(size_t)k * N] : 0.f; }
  __device__ void store(int, int m, int n, float acc) const { grad[(size_t)m * N + n] = acc; }
};

// ---------------------------------------------------------------- conv input gradient
// One GEMM per stride-parity class z = py*S + px: input pixels y = py + S*yy, x = px + S*xx
// only see taps kh = py + S*th, kw = px + S*tw, so K = (KS/S)^2 * COUT with no wasted taps.
// dx[b,y,x,c] = relu'(act_in) * sum_{th,tw,o} dy[b, yy-th, xx-tw, o] * W[kh,kw,c,o]
template <class L>
struct ConvDgrad {
  static constexpr const char* NAME = "conv_dgrad";
  static constexpr int BM = 64, BN = (L::CIN < 64 ? L::CIN : 64);
  static constexpr bool A_K_CONTIG = true, B_N_CONTIG = false;
  static constexpr int S = L::STRIDE, T = L::KS / L::STRIDE;
  static constexpr int NY = (L::HIN + S - 1) / S, NX = (L::WIN + S - 1) / S;
  static_assert(L::KS % L::STRIDE == 0, "taps per parity class must be uniform");
  int M, N, K;              // M = B*NY*NX, N = CIN, K = T*T*COUT
  const float* dy;          // [B,HOUT,WOUT,COUT], already relu-masked
  const float* w;           // HWIO
  const float* act_in;      // layer input (post-relu output of the previous layer)
  float* dx;
  struct ARow { int b, yy, xx; };
  struct BCol { const float* p; int py, px; };
  __device__ ARow a_row(int, int m) const {
    if (m >= M) return {-1, 0, 0};
    int b = m / (NY * NX), r = m - b * (NY * NX);
    int yy = r / NX;
    return {b, yy, r - yy * NX};
  }
  __device__ float a(const ARow& r, int k) const {
    if (r.b < 0 || k >= K) return 0.f;
    int tap = k / L::COUT, o = k - tap * L::COUT;
    int th = tap / T, tw = tap - th * T;
    int i = r.yy - th, j = r.xx - tw;
    if (i < 0 || i >= L::HOUT || j < 0 || j >= L::WOUT) return 0.f;
    return dy[((size_t)(r.b * L::HOUT + i) * L::WOUT + j) * L::COUT + o];
  }
  __device__ BCol b_col(int z, int n) const { return {n < N ? w + (size_t)n * L::COUT : nullptr, z / S, z % S}; }
  __device__ float b(const BCol& c, int k) const {
    if (!c.p || k >= K) return 0.f;
    int tap = k / L::COUT, o = k - tap * L::COUT;
    int th = tap / T, tw = tap - th * T;
    int kh = c.py + S * th, kw = c.px + S * tw;
    return c.p[(size_t)(kh * L::KS + kw) * (L::CIN * L::COUT) + o];
  }
  __device__ void store(int z, int m, int n, float acc) const {
    int b = m / (NY * NX), r = m - b * (NY * NX);
    int yy = r / NX, xx = r - yy * NX;
    int y = (z / S) + S * yy, x = (z % S) + S * xx;
    if (y >= L::HIN || x >= L::WIN) return;
    size_t idx = ((size_t)(b * L::HIN + y) * L::WIN + x) * L::CIN + n;
    dx[idx] = act_in[idx] > 0.f ? acc : 0.f;
  }
};

// ---------------------------------------------------------------- dense layers (fc1)
// y[m,n] = relu(sum_k x[m,k] W[k,n] + bias[n])
struct DenseFwd {
  static constexpr const char* NAME = "fc1_fwd";
  static constexpr int BM = 32, BN = 64;
  static constexpr bool A_K_CONTIG = true, B_N_CONTIG = true;
  int M, N, K;
  const float* x[kMaxNets];
  const float* w[kMaxNets];
  const float* bias[kMaxNets];
  float* y[kMaxNets];
  struct ARow { const float* p; };
  struct BCol { const float* p; };
  __device__ ARow a_row(int z, int m) const { return {m < M ? x[z] + (size_t)m * K : nullptr}; }
  __device__ float a(const ARow& r, int k) const { return (r.p && k < K) ? r.p[k] : 0.f; }
  __device__ BCol b_col(int z, int n) const { return {n < N ? w[z] + n : nullptr}; }
  __device__ float b(const BCol& c, int k) const { return (c.p && k < K) ? c.p[(size_t)k * N] : 0.f; }
  __device__ void store(int z, int m, int n, float acc) const {
    const float v = acc + bias[z][n];
    y[z][(size_t)m * N + n] = linear ? v : fmaxf(v, 0.f);
  }
  int linear = 0;
};

// dW[m'=k_in, n] = sum_b x[b,m'] * dy[b,n];  row m' == IN is ones -> bias gradient.
struct DenseWgrad {
  static constexpr const char* NAME = "fc1_wgrad";
  static constexpr int BM = 64, BN = 64;
  static constexpr bool A_K_CONTIG = false, B_N_CONTIG = true;
  int M, N, K;              // M = IN + 1, N = OUT, K = batch
  const float* x;           // [batch, IN]
  const float* dy;          // [batch, OUT]
  float* grad;
  struct ARow { int m; };
  struct BCol { const float* p; };
  __device__ ARow a_row(int, int m) const { return {m < M ? m : -1}; }
  __device__ float a(const ARow& r, int k) const {
    if (r.m < 0 || k >= K) return 0.f;
    return r.m == M - 1 ? 1.f : x[(size_t)k * (M - 1) + r.m];
  }
  __device__ BCol b_col(int, int n) const { return {n < N ? dy + n : nullptr}; }
  __device__ float b(const BCol& c, int k) const { return (c.p && k < K) ? c.p[(size_t)k * N] : 0.f; }
  __device__ void store(int, int m, int n, float acc) const { grad[(size_t)m * N + n] = acc; }
};

// dx[m,n=k_in] = relu'(act[m,n]) * sum_k dy[m,k] * W[n,k]
struct DenseDgrad {
  static constexpr const char* NAME = "fc1_dgrad";
  static constexpr int BM = 32, BN = 64;
  static constexpr bool A_K_CONTIG = true, B_N_CONTIG = false;
  int M, N, K;              // M = batch, N = IN, K = OUT
  const float* dy;          // [batch, OUT]
  const float* w;           // [IN, OUT]
  const float* act;         // [batch, IN] forward activation feeding this layer
  float* dx;
  struct ARow { const float* p; };
  struct BCol { const float* p; };
  __device__ ARow a_row(int, int m) const { return {m < M ? dy + (size_t)m * K : nullptr}; }
  __device__ float a(const ARow& r, int k) const { return (r.p && k < K) ? r.p[k] : 0.f; }
  __device__ BCol b_col(int, int n) const { return {n < N ? w + (size_t)n * K : nullptr}; }
  __device__ float b(const BCol& c, int k) const { return (c.p && k < K) ? c.p[k] : 0.f; }
  __device__ void store(int, int m, int n, float acc) const {
    size_t idx = (size_t)m * N + n;
    dx[idx] = act[idx] > 0.f ? acc : 0.f;
  }
};

}  // namespace dqnb
