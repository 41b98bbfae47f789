// tcgen05.mma issue-rate probe: one CTA per SM, one thread issues ITERS k-blocks of 4 (tf32, K=8) or
// 2 (f16, K=16) MMAs with M=128 against zeroed shared memory, for each operand-major combination and
// N in {64, 128, 256}; prints SM clocks per k-block (32 floats / 64 halves of K = 128 bytes per row).
//   nvcc -gencode arch=compute_100a,code=sm_100a -I dqn_mgsc_zoo_b200/csrc -o /tmp/mma_rate scripts/micro/mma_rate.cu
#include <cstdio>
#include <cstdlib>
#include "gemm_tf32.cuh"
using namespace dqnb;
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

__device__ __forceinline__ void f16_mma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, bool acc) {
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"((uint32_t)acc)
      : "memory");
}

// mode: 0 tf32 K/K, 1 tf32 K/MN, 2 tf32 MN/MN, 3 bf16 K/K, 4 tf32 K/K committing after every k-block
__global__ void __launch_bounds__(128, 1) probe(int n, int mode, int iters, long long* out) {
  extern __shared__ uint8_t raw[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const uint32_t smem = (smem_u32(raw) + 1023u) & ~1023u;
  for (int i = threadIdx.x; i < 48 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(raw)[i] = 0u;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  if (threadIdx.x < 32) t32_alloc<256>(&slot);
  fence_proxy_async_smem();
  t32_fence_before();
  __syncthreads();
  t32_fence_after();
  const uint32_t tmem = slot;
  if (threadIdx.x == 0) {
    const bool a_mn = mode == 2, b_mn = mode == 1 || mode == 2, f16 = mode == 3;
    uint32_t idesc = t32_idesc(n, a_mn, b_mn);
    if (f16) idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);   // bf16 x bf16 -> f32
    const uint64_t ad = a_mn ? t32_desc(smem, 4096, 512, 1) : t32_desc(smem, 16, 1024);
    const uint64_t bd = b_mn ? t32_desc(smem + 16384, 4096, 512, 1) : t32_desc(smem + 16384, 16, 1024);
    const uint32_t aks = (a_mn ? 1024 : 32) >> 4, bks = (b_mn ? 1024 : 32) >> 4;
    uint32_t ph = 0;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      if (f16) {
        for (int ks = 0; ks < 4; ++ks) f16_mma(tmem, ad + ks * 2, bd + ks * 2, idesc, true);   // 4 x K=16 halves = 128 B rows
      } else {
        for (int ks = 0; ks < 4; ++ks) t32_mma(tmem, ad + ks * aks, bd + ks * bks, idesc, true);
      }
      if (mode == 4) { t32_commit(&bar); mbar_wait(&bar, ph); ph ^= 1; }
    }
    if (mode != 4) { t32_commit(&bar); mbar_wait(&bar, 0); }
    const long long t1 = clock64();
    if (blockIdx.x == 0) *out = t1 - t0;
  }
  __syncthreads();
  if (threadIdx.x < 32) t32_dealloc<256>(tmem);
}

int main() {
  long long* d;
  CK(cudaMalloc(&d, 8));
  CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  const char* names[5] = {"tf32 A:K  B:K ", "tf32 A:K  B:MN", "tf32 A:MN B:MN", "bf16 A:K  B:K ", "tf32 K/K commit+wait per k-block"};
  const int iters = 2000;
  for (int mode = 0; mode < 5; ++mode)
    for (int n : {64, 128, 256}) {
      long long h = 0;
      for (int rep = 0; rep < 2; ++rep) {
        probe<<<148, 128, 64 * 1024>>>(n, mode, iters, d);
        CK(cudaDeviceSynchronize());
      }
      CK(cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost));
      const double clk = (double)h / iters;
      const double flop = 2.0 * 128 * n * (mode == 3 ? 64 : 32);
      printf("%s N=%3d: %7.1f clk per k-block  -> %6.0f FLOP/clk/SM = %6.0f TFLOP/s at 1.965 GHz x 148\n", names[mode], n, clk,
             flop / clk, flop / clk * 1.965e9 * 148 / 1e12);
    }
  return 0;
}
