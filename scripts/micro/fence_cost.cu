// Microbenchmark: what does fence.proxy.async.shared::cta (MEMBAR.ALL.CTA + FENCE.VIEW.ASYNC.S) cost
// while the same thread has a global load / a cp.async in flight?   nvcc -arch=sm_100a fence_cost.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
// mode 0: fence only; 1: LDG in flight (result unused until after); 2: cp.async in flight (not waited);
// 3: LDG in flight, no fence (baseline for the timing code)
__global__ void k(int mode, const uint4* src, size_t stride, long long* out, uint4* sink) {
  __shared__ uint4 buf[256];
  __shared__ float st[256];
  const int t = threadIdx.x;
  long long tot = 0;
  uint4 acc = make_uint4(0, 0, 0, 0);
  for (int it = 0; it < 64; ++it) {
    const uint4* p = src + ((size_t)(blockIdx.x * 64 + it) * 256 + t) * stride;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (mode == 1 || mode == 3) v = __ldcg(p);
    if (mode == 2) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(&buf[t])), "l"(p) : "memory");
    st[t] = (float)it;
    long long t0 = clock64();
    if (mode != 3) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    long long t1 = clock64();
    tot += t1 - t0;
    acc.x += v.x; acc.y += v.y;
    if (mode == 2) asm volatile("cp.async.wait_all;" ::: "memory");
    __syncwarp();
  }
  if (t == 0) out[blockIdx.x] = tot / 64;
  if (acc.x == 0x12345) sink[0] = acc;
}
int main() {
  size_t n = (size_t)148 * 64 * 256 * 8;
  uint4* src; cudaMalloc(&src, n * sizeof(uint4)); cudaMemset(src, 1, n * sizeof(uint4));
  long long* out; cudaMalloc(&out, 148 * 8); uint4* sink; cudaMalloc(&sink, 64);
  for (int mode = 0; mode < 4; ++mode) {
    for (int rep = 0; rep < 2; ++rep) k<<<148, 256>>>(mode, src, 8, out, sink);
    cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
    long long s = 0; for (int i = 0; i < 148; ++i) s += h[i];
    printf("mode %d: avg cycles around fence = %lld (%s)\n", mode, s / 148, cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
