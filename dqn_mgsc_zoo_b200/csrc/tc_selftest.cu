// Test-only entry point: runs gemm_tc on plain fp32 matrices so the tcgen05 path (descriptors,
// swizzled staging, TMEM epilogue, split-K fix-up) can be checked against NumPy on the GPU box.
#include "gemm_tc.cuh"
#include "gemm_tf32.cuh"

using namespace dqnb;

namespace {
template <int BN_, bool A_MN, bool B_MN>
struct PlainGemm {
  static constexpr int BN = BN_;
  static constexpr bool A_MN_MAJOR = A_MN, B_MN_MAJOR = B_MN;
  int M, N, K;
  const float* a;   // K-major: [M,K] row-major; MN-major: [K,M] row-major
  const float* b;   // K-major: [N,K] row-major; MN-major: [K,N] row-major
  float* c;         // [M,N]
  __device__ static uint4 load8(const float* base, int rows, int cols, int r, int c0) {
    // 8 consecutive elements base[r][c0..c0+7] of a [rows, cols] row-major matrix, zero-filled
    float f[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) f[i] = (r < rows && c0 + i < cols) ? base[(size_t)r * cols + c0 + i] : 0.f;
    return pack8_bf16(f);
  }
  __device__ uint4 a_chunk(int, int i, int j) const { return A_MN ? load8(a, K, M, j, i) : load8(a, M, K, i, j); }
  __device__ uint4 b_chunk(int, int i, int j) const { return B_MN ? load8(b, K, N, j, i) : load8(b, N, K, i, j); }
  __device__ void store_row(int, int m, int n0, const float (&v)[16], int valid) const {
    for (int i = 0; i < valid; ++i) c[(size_t)m * N + n0 + i] = v[i];
  }
};

template <class P>
int run(const float* a, const float* b, float* c, int M, int N, int K, int splits, cudaStream_t st) {
  P p;
  p.M = M; p.N = N; p.K = K; p.a = a; p.b = b; p.c = c;
  SplitK sk{nullptr, nullptr};
  if (splits > 1) {
    DQNB_CUDA(cudaMalloc(&sk.ws, sizeof(float) * gemm_tc_ws_floats(p, 1, splits)));
    DQNB_CUDA(cudaMalloc(&sk.counters, sizeof(unsigned) * gemm_tc_tiles(p, 1)));
    DQNB_CUDA(cudaMemsetAsync(sk.counters, 0, sizeof(unsigned) * gemm_tc_tiles(p, 1), st));
  }
  DQNB_CUDA(launch_gemm_tc(p, 1, splits, sk, st));
  DQNB_CUDA(cudaStreamSynchronize(st));
  if (sk.ws) { cudaFree(sk.ws); cudaFree(sk.counters); }
  return DQNB_OK;
}
}  // namespace

// C[M,N] = A * B^T with bf16-rounded operands; device pointers; bn in {32, 64, 128}.
extern "C" int dqnb_selftest_tc_gemm(const float* a_dev, const float* b_dev, float* c_dev, int M, int N, int K,
                                     int a_mn_major, int b_mn_major, int bn, int splits, uintptr_t stream) {
  DQNB_REQUIRE(a_dev && b_dev && c_dev && M > 0 && N > 0 && K > 0 && splits >= 1, "bad argument");
  cudaStream_t st = as_stream(stream);
#define DQNB_CASE(BN, AM, BM) \
  if (bn == BN && a_mn_major == AM && b_mn_major == BM) return run<PlainGemm<BN, AM, BM>>(a_dev, b_dev, c_dev, M, N, K, splits, st);
  DQNB_CASE(32, 0, 0) DQNB_CASE(64, 0, 0) DQNB_CASE(128, 0, 0)
  DQNB_CASE(64, 1, 0) DQNB_CASE(64, 0, 1) DQNB_CASE(64, 1, 1) DQNB_CASE(32, 0, 1) DQNB_CASE(128, 1, 1)
#undef DQNB_CASE
  set_error("selftest combination bn=%d a_mn=%d b_mn=%d not instantiated", bn, a_mn_major, b_mn_major);
  return DQNB_EINVAL;
}

// ---- TF32 slab kernel (gemm_tf32.cuh) on plain fp32 matrices, all four operand-major combinations.
namespace {
template <int BN_, bool A_MN, bool B_MN>
struct PlainGemm32 {
  static constexpr int BN = BN_, KB_MAX = BN_ > 64 ? 4 : 8;
  static constexpr bool A_MN_MAJOR = A_MN, B_MN_MAJOR = B_MN, DEFER_REDUCE = false;
  int M, N, K;
  const float* a;   // K-major: [M,K] row-major; MN-major: [K,M] row-major
  const float* b;   // K-major: [N,K] row-major; MN-major: [K,N] row-major
  float* c;         // [M,N]
  __device__ static void chunk(const float* base, int rows, int cols, bool mn, int i, int j, uint32_t dst) {
    // K-major: base[i][j..j+3] of [rows, cols];  MN-major: base[j][i..i+3] of [cols(k), rows(mn)]
    bool ok = i < rows && j < cols;
    const float* src = mn ? base + (size_t)j * rows + i : base + (size_t)i * cols + j;
    cp16(dst, ok ? src : base, ok);
  }
  __device__ void a_chunk(int, int i, int j, uint32_t dst) const { chunk(a, M, K, A_MN, i, j, dst); }
  __device__ void b_chunk(int, int i, int j, uint32_t dst) const { chunk(b, N, K, B_MN, i, j, dst); }
  __device__ void store_row(int, int m, int n0, const float (&v)[16], int valid) const {
    for (int i = 0; i < valid; ++i) c[(size_t)m * N + n0 + i] = v[i];
  }
  __device__ void store_partial(int, int, int, int, const float (&)[16], int) const {}
};

template <class P>
int run32(const float* a, const float* b, float* c, int M, int N, int K, int splits, cudaStream_t st) {
  P p;
  p.M = M; p.N = N; p.K = K; p.a = a; p.b = b; p.c = c;
  SplitK sk{nullptr, nullptr};
  if (splits > 1) {
    DQNB_CUDA(cudaMalloc(&sk.ws, sizeof(float) * gemm_tf32_ws_floats(p, 1, splits)));
    DQNB_CUDA(cudaMalloc(&sk.counters, sizeof(unsigned) * gemm_tf32_tiles(p, 1)));
    DQNB_CUDA(cudaMemsetAsync(sk.counters, 0, sizeof(unsigned) * gemm_tf32_tiles(p, 1), st));
  }
  DQNB_CUDA(launch_gemm_tf32(p, 1, splits, sk, st));
  DQNB_CUDA(cudaStreamSynchronize(st));
  if (sk.ws) { cudaFree(sk.ws); cudaFree(sk.counters); }
  return DQNB_OK;
}
}  // namespace

// C[M,N] = A * B^T on the TF32 tensor-core path; M, N, K multiples of 4; splits >= ceil(K/256)
// (ceil(K/128) for bn=128).
extern "C" int dqnb_selftest_tf32_gemm(const float* a_dev, const float* b_dev, float* c_dev, int M, int N, int K,
                                       int a_mn_major, int b_mn_major, int bn, int splits, uintptr_t stream) {
  DQNB_REQUIRE(a_dev && b_dev && c_dev && M > 0 && N > 0 && K > 0 && splits >= 1, "bad argument");
  DQNB_REQUIRE(M % 4 == 0 && N % 4 == 0 && K % 4 == 0, "M, N, K must be multiples of 4");
  cudaStream_t st = as_stream(stream);
#define DQNB_CASE(BN, AM, BM) \
  if (bn == BN && a_mn_major == AM && b_mn_major == BM) return run32<PlainGemm32<BN, AM, BM>>(a_dev, b_dev, c_dev, M, N, K, splits, st);
  DQNB_CASE(32, 0, 0) DQNB_CASE(64, 0, 0) DQNB_CASE(128, 0, 0) DQNB_CASE(64, 1, 0) DQNB_CASE(64, 0, 1)
  DQNB_CASE(64, 1, 1) DQNB_CASE(32, 0, 1) DQNB_CASE(32, 1, 1) DQNB_CASE(128, 1, 1) DQNB_CASE(128, 0, 1)
#undef DQNB_CASE
  set_error("selftest combination bn=%d a_mn=%d b_mn=%d not instantiated", bn, a_mn_major, b_mn_major);
  return DQNB_EINVAL;
}
