// Acting forward (select_action, dqn/agent.py:121-131): q(s) of ONE observation with one net, as a
// single persistent kernel.  The layer-by-layer GEMM kernels cost ~7 us each whatever the batch, so a
// batch-1 forward was a 40 us chain of five launches; here 148 CTAs walk the five layers with a
// grid-wide barrier between them (fp32 FMA, weights straight from the fp32 arena, fixed summation
// order: deterministic).  One warp per task:
//   conv1/2/3: one CTA per output position (400 / 81 / 49): patch in smem, K split over the 8 warps,
//          lanes = channels, every warp's weight loads issued up front, slice sums added in warp order
//   fc1  : (32 outputs, 64-wide K slice) 16 x 49 tasks -> partials[49][512]
//   fc2  : CTA 0: h1 = relu(bias + sum of the 49 partials), then one warp per action dots it with its weight column
// Layouts as in nature_problems.cuh: activations NHWC, weights HWIO = row-major [K][N].
#pragma once
#include "nature_problems.cuh"

namespace dqnb {

struct GridBarrier {
  unsigned count, gen;
};
// Self-resetting generation barrier (state persists across launches).  All CTAs of the grid must be
// co-resident: grid = one CTA per SM, checked with the occupancy API, and the kernel is launched with
// cudaLaunchAttributeCooperative so the grid is scheduled all-or-nothing.
__device__ __forceinline__ void grid_barrier(GridBarrier* b) {
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned gen = *reinterpret_cast<volatile unsigned*>(&b->gen);
    __threadfence();
    if (atomicAdd(&b->count, 1u) == gridDim.x - 1) {
      b->count = 0;
      __threadfence();
      atomicAdd(&b->gen, 1u);
    } else {
      while (*reinterpret_cast<volatile unsigned*>(&b->gen) == gen) {
      }
    }
    __threadfence();
  }
  __syncthreads();
}

struct ActArgs {
  const uint8_t* obs;        // [84][84][4]
  const float *w1, *b1, *w2, *b2, *w3, *b3, *fw1, *fb1, *fw2, *fb2;
  float *a1, *a2, *a3;       // [400*32], [81*64], [49*64]
  float* part;               // [49][512] fc1 partial sums
  float* q;                  // [A]
  float* q_host;             // optional pinned copy of q (null = off)
  unsigned* seq_dev;         // launches executed so far (device counter)
  volatile unsigned* done_host;   // pinned, device-mapped: receives the launch's sequence number after q_host is written
  GridBarrier* bar;
  int num_actions, shared_bias;
};

constexpr int kActThreads = 256, kActWarps = kActThreads / 32;
constexpr int kActSlice = 64, kActSlices = kFc1In / kActSlice;   // 49
static_assert(kFc1In % kActSlice == 0, "fc1 K slices");

// One conv layer, one CTA per output position: the patch is staged in shared memory by the whole CTA,
// warp w owns the K slice [w K/8, (w+1) K/8) and lane l the channels l, l+32, ...; all weight loads of a
// slice are issued before the FMAs (the filters come from L2 - few positions per SM, no L1 reuse - so
// load depth, not FLOPs, sets the time), then the 8 slice sums are added in warp order.
template <class L, class TIn>
__device__ __forceinline__ void act_conv_layer(const TIn* __restrict__ x, const float* __restrict__ w,
                                               const float* __restrict__ bias, float* __restrict__ y, float* patch,
                                               float* partial, float scale) {
  constexpr int RUN = L::ROWLEN, KW = L::KCOL / kActWarps, NC = L::COUT / 32, HALF = KW / 2;
  static_assert(L::KCOL % (2 * kActWarps) == 0 && L::COUT % 32 == 0, "K slices / channel groups");
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int p = blockIdx.x; p < L::HOUT * L::WOUT; p += gridDim.x) {
    const int i = p / L::WOUT, j = p - i * L::WOUT;
    __syncthreads();   // previous position's patch / partials are dead
    for (int k = threadIdx.x; k < L::KCOL; k += kActThreads) {
      const int r = k / RUN, o = k - r * RUN;
      const TIn* src = x + ((size_t)(i * L::STRIDE + r) * L::WIN + j * L::STRIDE) * L::CIN + o;
      patch[k] = sizeof(TIn) == 1 ? (float)*src : __ldcg(reinterpret_cast<const float*>(src));
    }
    __syncthreads();
    float acc[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) acc[c] = 0.f;
    const float* wk = w + (size_t)(warp * KW) * L::COUT + lane;
    const float* pk = patch + warp * KW;
#pragma unroll
    for (int hblk = 0; hblk < 2; ++hblk) {
      float wv[HALF][NC];
#pragma unroll
      for (int k = 0; k < HALF; ++k)
#pragma unroll
        for (int c = 0; c < NC; ++c) wv[k][c] = __ldg(wk + (size_t)(hblk * HALF + k) * L::COUT + 32 * c);
#pragma unroll
      for (int k = 0; k < HALF; ++k)
#pragma unroll
        for (int c = 0; c < NC; ++c) acc[c] = fmaf(pk[hblk * HALF + k], wv[k][c], acc[c]);
    }
#pragma unroll
    for (int c = 0; c < NC; ++c) partial[warp * L::COUT + 32 * c + lane] = acc[c];
    __syncthreads();
    if (threadIdx.x < L::COUT) {
      float v = 0.f;
#pragma unroll
      for (int q = 0; q < kActWarps; ++q) v += partial[q * L::COUT + threadIdx.x];
      y[(size_t)p * L::COUT + threadIdx.x] = fmaxf(v * scale + __ldg(bias + threadIdx.x), 0.f);
    }
  }
}

__global__ void __launch_bounds__(kActThreads, 1) act_forward_kernel(ActArgs g) {
  __shared__ float patch_s[Conv3::KCOL];            // 576 floats: the largest patch; later h1
  __shared__ float partial_s[kActWarps * 64];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int warp_global = blockIdx.x * kActWarps + warp, warps_total = gridDim.x * kActWarps;

  // conv1 reads the uint8 frames; the 1/255 of networks.py:193 is applied to the sum
  act_conv_layer<Conv1, uint8_t>(g.obs, g.w1, g.b1, g.a1, patch_s, partial_s, 1.0f / 255.0f);
  grid_barrier(g.bar);
  act_conv_layer<Conv2, float>(g.a1, g.w2, g.b2, g.a2, patch_s, partial_s, 1.0f);
  grid_barrier(g.bar);
  act_conv_layer<Conv3, float>(g.a2, g.w3, g.b3, g.a3, patch_s, partial_s, 1.0f);
  grid_barrier(g.bar);

  // ---- fc1 partials: task = (group of 32 outputs, K slice)
  for (int task = warp_global; task < (kFc1Out / 32) * kActSlices; task += warps_total) {
    const int s = task / (kFc1Out / 32), n = (task - s * (kFc1Out / 32)) * 32 + lane;
    const float* xs = g.a3 + s * kActSlice;
    const float x0 = __ldcg(xs + lane), x1 = __ldcg(xs + 32 + lane);
    const float* wc = g.fw1 + (size_t)(s * kActSlice) * kFc1Out + n;
    // the 6.4 MB of fc1 weights come from L2 / HBM: all 64 loads of the slice are issued before the FMAs
    float wv[kActSlice];
#pragma unroll
    for (int k = 0; k < kActSlice; ++k) wv[k] = __ldg(wc + (size_t)k * kFc1Out);
    float acc0 = 0.f, acc1 = 0.f;
#pragma unroll
    for (int k = 0; k < 32; ++k) {
      acc0 = fmaf(__shfl_sync(0xffffffffu, x0, k), wv[k], acc0);
      acc1 = fmaf(__shfl_sync(0xffffffffu, x1, k), wv[k + 32], acc1);
    }
    g.part[s * kFc1Out + n] = acc0 + acc1;
  }
  grid_barrier(g.bar);

  // ---- fc2 (CTA 0): h1 = relu(bias + the 49 partials in slice order), two entries per thread with all
  // loads in flight at once; then one warp per action dots h1 with its weight column.
  if (blockIdx.x != 0) return;
  float* h1_s = patch_s;   // 512 floats (the patches are dead)
  static_assert(kFc1Out == 2 * kActThreads && Conv3::KCOL >= kFc1Out, "h1 staging");
  {
    float pv[2][kActSlices];
#pragma unroll
    for (int u = 0; u < 2; ++u)
#pragma unroll
      for (int sl = 0; sl < kActSlices; ++sl) pv[u][sl] = __ldcg(g.part + sl * kFc1Out + threadIdx.x + u * kActThreads);
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      float h = __ldg(g.fb1 + threadIdx.x + u * kActThreads);
#pragma unroll
      for (int sl = 0; sl < kActSlices; ++sl) h += pv[u][sl];
      h1_s[threadIdx.x + u * kActThreads] = fmaxf(h, 0.f);
    }
  }
  __syncthreads();
  const int A = g.num_actions;
  for (int a = warp; a < A; a += kActWarps) {
    float acc = 0.f;
#pragma unroll
    for (int k = lane; k < kFc1Out; k += 32) acc = fmaf(h1_s[k], __ldg(g.fw2 + (size_t)k * A + a), acc);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) {
      const float v = acc + __ldg(g.fb2 + (g.shared_bias ? 0 : a));
      g.q[a] = v;
      if (g.q_host != nullptr) g.q_host[a] = v;
    }
  }
  // completion flag for the host (it polls pinned memory instead of synchronising the stream)
  if (g.done_host != nullptr) {
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
      const unsigned seq = *g.seq_dev + 1;
      *g.seq_dev = seq;
      *g.done_host = seq;
    }
  }
}

}  // namespace dqnb
