// Test-only entry point: runs the TF32 slab GEMM on plain fp32 matrices so the tcgen05 path (descriptors,
// swizzled staging, TMEM epilogue, cluster / deferred split-K) can be checked against NumPy on the GPU box.
#include "gemm_tf32.cuh"

using namespace dqnb;

// ---- TF32 slab kernel (gemm_tf32.cuh) on plain fp32 matrices, all four operand-major combinations.
namespace {
// DEFER: split-K partials go to part[split][M][N] and sum_partials_kernel adds them (the learner's
// weight-gradient / fc1 scheme); otherwise the splits form a cluster and reduce over DSMEM.
template <int BN_, bool A_MN, bool B_MN, bool DEFER>
struct PlainGemm32 {
  static constexpr int BN = BN_, KB_MAX = BN_ > 64 ? 4 : 8;
  static constexpr bool A_MN_MAJOR = A_MN, B_MN_MAJOR = B_MN, DEFER_REDUCE = DEFER;
  float* part;
  int M, N, K;
  const float* a;   // K-major: [M,K] row-major; MN-major: [K,M] row-major
  const float* b;   // K-major: [N,K] row-major; MN-major: [K,N] row-major
  float* c;         // [M,N]
  // contexts: K-major row = {&base[i][0], 0}, col = {-, j};  MN-major row = {-, i}, col = {&base[j][0], 0}
  struct Ctx { const float* p; int o; };
  using ARow = Ctx; using ACol = Ctx; using BRow = Ctx; using BCol = Ctx;
  __device__ static Ctx mk_row(const float* base, int rows, int cols, bool mn, int i) {
    if (mn) return {nullptr, i < rows ? i : -1};
    return {i < rows ? base + (size_t)i * cols : nullptr, 0};
  }
  __device__ static Ctx mk_col(const float* base, int rows, int cols, bool mn, int j) {
    if (mn) return {j < cols ? base + (size_t)j * rows : nullptr, 0};
    return {nullptr, j < cols ? j : -1};
  }
  __device__ static void put(const float* base, bool mn, const Ctx& r, const Ctx& c, uint32_t dst) {
    const float* ptr = mn ? c.p : r.p;
    const int off = mn ? r.o : c.o;
    const bool ok = ptr != nullptr && off >= 0;
    cp16(dst, ok ? ptr + off : base, ok);
  }
  __device__ ARow a_row(int, int i) const { return mk_row(a, M, K, A_MN, i); }
  __device__ ACol a_col(int, int j) const { return mk_col(a, M, K, A_MN, j); }
  __device__ void a_put(const ARow& r, const ACol& c, uint32_t dst) const { put(a, A_MN, r, c, dst); }
  __device__ BRow b_row(int, int i) const { return mk_row(b, N, K, B_MN, i); }
  __device__ BCol b_col(int, int j) const { return mk_col(b, N, K, B_MN, j); }
  __device__ void b_put(const BRow& r, const BCol& c, uint32_t dst) const { put(b, B_MN, r, c, dst); }
  __device__ void store4(int, int split, int m, int n, float4 v) const {
    float* dst = DEFER ? part + (size_t)split * M * N : c;
    *reinterpret_cast<float4*>(dst + (size_t)m * N + n) = v;
  }
};
__global__ void sum_partials_kernel(const float* part, float* c, int64_t n, int splits) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float acc = 0.f;
  for (int q = 0; q < splits; ++q) acc += part[(size_t)q * n + i];
  c[i] = acc;
}

template <class P>
int run32(const float* a, const float* b, float* c, int M, int N, int K, int splits, cudaStream_t st) {
  P p;
  p.M = M; p.N = N; p.K = K; p.a = a; p.b = b; p.c = c; p.part = nullptr;
  if (P::DEFER_REDUCE) DQNB_CUDA(cudaMalloc(&p.part, sizeof(float) * (size_t)splits * M * N));
  SplitK sk{};
  cudaError_t e = launch_gemm_tf32(p, 1, splits, sk, st);
  if (e == cudaSuccess && P::DEFER_REDUCE) {
    const int64_t n = (int64_t)M * N;
    sum_partials_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(p.part, c, n, splits);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (p.part) cudaFree(p.part);
  DQNB_CUDA(e);
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
  if (bn == BN && a_mn_major == AM && b_mn_major == BM) {                                                         \
    if (splits > kT32MaxCluster) return run32<PlainGemm32<BN, AM, BM, true>>(a_dev, b_dev, c_dev, M, N, K, splits, st); \
    return run32<PlainGemm32<BN, AM, BM, false>>(a_dev, b_dev, c_dev, M, N, K, splits, st);                         \
  }
  DQNB_CASE(32, 0, 0) DQNB_CASE(64, 0, 0) DQNB_CASE(128, 0, 0) DQNB_CASE(64, 1, 0) DQNB_CASE(64, 0, 1)
  DQNB_CASE(64, 1, 1) DQNB_CASE(32, 0, 1) DQNB_CASE(32, 1, 1) DQNB_CASE(128, 1, 1) DQNB_CASE(128, 0, 1)
#undef DQNB_CASE
  set_error("selftest combination bn=%d a_mn=%d b_mn=%d not instantiated", bn, a_mn_major, b_mn_major);
  return DQNB_EINVAL;
}
