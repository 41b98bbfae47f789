// Shared helpers for libdqnb200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <string>

#include "../../include/dqnb200.h"

namespace dqnb {

void set_error(const char* fmt, ...);

#define DQNB_CUDA(expr)                                                                   \
  do {                                                                                    \
    cudaError_t _e = (expr);                                                              \
    if (_e != cudaSuccess) {                                                              \
      dqnb::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
      return DQNB_ECUDA;                                                                  \
    }                                                                                     \
  } while (0)

#define DQNB_REQUIRE(cond, ...)          \
  do {                                   \
    if (!(cond)) {                       \
      dqnb::set_error(__VA_ARGS__);      \
      return DQNB_EINVAL;                \
    }                                    \
  } while (0)

#define DQNB_TRY(expr)           \
  do {                           \
    int _rc = (expr);            \
    if (_rc != DQNB_OK) return _rc; \
  } while (0)

struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    cudaGetDevice(&prev);
    if (prev != dev) cudaSetDevice(dev);
    else prev = -1;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

inline cudaStream_t as_stream(uintptr_t s) { return reinterpret_cast<cudaStream_t>(s); }

constexpr int kNumSMs = 148;  // B200

// ---- device-side PTX wrappers (mbarrier + 1-D bulk async copy, i.e. the TMA engine) ----
#ifdef __CUDACC__
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// One arrival per warp instead of 32: every lane has already fenced its own writes; the warp
// converges (which orders those fences before lane 0's release-arrive) and lane 0 arrives.  256
// per-thread arrivals on one shared-memory barrier serialise (~3 clk each, measured as ~860 clk per
// ring stage in the stream kernel); barriers fed this way are initialised with the WARP count.
__device__ __forceinline__ void mbar_arrive_warp(uint64_t* bar) {
  __syncwarp();
  if ((threadIdx.x & 31) == 0) mbar_arrive(bar);
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// The same operations on a precomputed shared-space address: smem_u32() of a __shared__ object is
// S2R SR_CgaCtaId + LEA every time it is evaluated, which sits on the serial path of ring loops.
__device__ __forceinline__ void mbar_arrive_a(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_warp_a(uint32_t bar) {
  __syncwarp();
  if ((threadIdx.x & 31) == 0) mbar_arrive_a(bar);
}
__device__ __forceinline__ void mbar_wait_a(uint32_t bar, uint32_t parity) {
  for (uint32_t spins = 0;; ++spins) {
    uint32_t ok;
    asm volatile(
        "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    if (ok) return;
    if (spins > (1u << 24)) __trap();
  }
}
// Bounded wait: a barrier that never completes (a bug) traps instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  for (uint32_t spins = 0; !mbar_try_wait(bar, parity); ++spins)
    if (spins > (1u << 24)) __trap();
}
// global -> shared bulk copy; bytes and both addresses must be multiples of 16.
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes,
                                         uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(smem_u32(smem_dst)),
      "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
// Tiled TMA load through a tensor map (cuTensorMapEncodeTiled on the host; the 128-byte descriptor may live in
// global memory as long as the device never writes it): the box at coordinates (c0 = innermost, c1) lands at
// `smem_dst` in the map's swizzle mode and completes `bytes of the box` on the mbarrier.
__device__ __forceinline__ void tma_load_2d(uint32_t smem_dst, const void* tmap, int c0, int c1, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::
          "r"(smem_dst),
      "l"(tmap), "r"(c0), "r"(c1), "r"(bar)
      : "memory");
}
__device__ __forceinline__ void tma_prefetch_map(const void* tmap) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}
// shared -> global bulk copy (bulk async-group completion).
__device__ __forceinline__ void bulk_s2g(void* gmem_dst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst),
               "r"(smem_u32(smem_src)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
// Timeline stamps (diagnostics): every CTA folds its begin / end %globaltimer (ns) into the
// launch's slot: tl[0] = min begin, tl[1] = max end, tl[2] = max begin (launch ramp),
// tl[3] = longest single-CTA duration.  tl == nullptr: off.
constexpr int kTlSlot = 4;
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ unsigned long long tl_begin(unsigned long long* tl) {
  if (tl == nullptr || threadIdx.x != 0) return 0;
  const unsigned long long now = global_ns();
  atomicMin(tl, now);
  atomicMax(tl + 2, now);
  return now;
}
__device__ __forceinline__ void tl_end(unsigned long long* tl, unsigned long long began = 0) {
  if (tl == nullptr || threadIdx.x != 0) return;
  const unsigned long long now = global_ns();
  atomicMax(tl + 1, now);
  if (began) atomicMax(tl + 3, now - began);
}
// One lane of a converged warp (elect.sync): ptxas keeps what the elected lane computes from warp-uniform inputs in
// uniform registers; a `threadIdx.x == k` test makes it move every tcgen05 / TMA operand through a
// R2UR.BROADCAST + BRA.U.ANY waterfall loop (measured: 18 instructions per tcgen05.mma).
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n.reg .pred p;\nelect.sync _|p, 0xffffffff;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
#endif

// Round-to-nearest fp32 -> tf32 (10 explicit mantissa bits, low 13 bits cleared).  tcgen05's
// kind::tf32 simply ignores the low bits (truncation, a bias toward zero that adds up layer by
// layer), so every tensor that feeds a TF32 GEMM is rounded here by the kernel that produces it.
__device__ __forceinline__ float tf32_rn(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
__device__ __forceinline__ float4 tf32_rn4(float4 v) {
  return make_float4(tf32_rn(v.x), tf32_rn(v.y), tf32_rn(v.z), tf32_rn(v.w));
}

}  // namespace dqnb
