// K4 (TD / loss, warp-shuffle reductions), fc2 forward+backward, K5 fused centered RMSProp.
#pragma once
#include <cuda_fp16.h>
#include "nature_problems.cuh"

namespace dqnb {

constexpr int kMaxActions = 32;   // one lane per action in the warp-level argmax

struct HeadArgs {
  int batch, num_actions, shared_bias, double_q;
  float grad_bound;        // rlax.clip_gradient bound (dqn/agent.py:101-103)
  float inv_mean;          // 1 / (number of rows the loss mean runs over)
  const float* h1[kMaxNets];   // [batch,512] relu(fc1): online(s_tm1), target(s_t), online(s_t)
  // TF32 mode: fc1 leaves split-K partials [net][split][batch][512]; the head sums them in split
  // order, adds the bias, applies the relu and writes h1 (fc1_part == nullptr: h1 is read as is)
  const float* fc1_part;
  int fc1_splits;
  const float* b1[kMaxNets];   // fc1 bias
  float* h1_out[kMaxNets];
  const float* w2[kMaxNets];   // [512,A]
  const float* b2[kMaxNets];   // [A] or [1]
  const int32_t* action;
  const float* reward;
  const float* discount;
  const float* row_weight;  // optional per-row weight on the cotangent (MGSC), may be null
  int weight_inside;        // 1: the weight is an importance-sampling weight of the loss (prioritized replay): it
                            // multiplies the cotangent before the clip and the row loss
  float* q[kMaxNets];       // [batch,A] outputs
  float* td;                // [batch]
  float* row_loss;          // [batch] 0.5*td^2
  float* dq_taken;          // [batch] d loss / d q_tm1[b, a_b]
  float* dh1;               // [batch,512] gradient wrt fc1 pre-activation (relu-masked)
  unsigned long long* tl;   // optional timeline slot
  int round_tf32;           // dh1 feeds TF32 GEMMs: round it to nearest here
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// lane a holds q[a] (lanes >= A hold -inf): returns max value and its FIRST index (jnp.argmax).
__device__ __forceinline__ void warp_argmax(float v, int lane, float& best, int& best_idx) {
  best = v;
  best_idx = lane;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float ov = __shfl_xor_sync(0xffffffffu, best, o);
    int oi = __shfl_xor_sync(0xffffffffu, best_idx, o);
    if (ov > best || (ov == best && oi < best_idx)) { best = ov; best_idx = oi; }
  }
}

// One CTA per batch row, one group of 128 threads (4 warps) per net.  (1) h1 of the group's net (finishing fc1 when
// it was left as split-K partials: thread c of the group owns columns 4c..4c+3), (2) fc2: the group's (net, action)
// dot products, (3) warp 0 forms the q_learning / double_q_learning target (rlax), the td error, the clipped
// cotangent and the per-row loss, (4) all threads write the fc2 input gradient dh1.
// The kernel is a chain of dependent L2 round trips, so everything that does not depend on the preceding kernel
// (fc1 bias, fc2 weights and bias, the row's action / reward / discount / weight, the fc2 column of the taken
// action) is fetched BEFORE griddepcontrol.wait - under programmatic dependent launch that is while fc1 is still
// storing its partials - and all of a thread's partials are in flight at once.
// kDepth = split-K partials a thread keeps in flight: 20 for the small-batch slab path (17-18 splits, one CTA per SM
// is all a batch of 32 rows needs), 4 for large batches (1-3 splits; two CTAs per SM, the grid is thousands of rows).
constexpr int kHeadGroup = 128;
constexpr int kHeadThreads = kMaxNets * kHeadGroup;
template <int kDepth>
__global__ void __launch_bounds__(kHeadThreads, kDepth > 8 ? 1 : 2) head_td_kernel(HeadArgs g) {
  __shared__ __align__(16) float h_s[kFc1Out];
  __shared__ float q_s[kMaxNets][kMaxActions];
  __shared__ float red_s[kMaxNets][kHeadGroup / 32][kMaxActions];
  __shared__ float dq_s;
  const unsigned long long tl_t0 = tl_begin(g.tl);
  const int t = threadIdx.x, lane = t & 31;
  const int z = t / kHeadGroup, c = t - z * kHeadGroup, gw = c >> 5;   // net, column group, warp inside the net's group
  const int row = blockIdx.x;
  const int A = g.num_actions, nets = g.double_q ? 3 : 2;
  const bool on = z < nets;
  static_assert(kFc1Out == 4 * kHeadGroup, "one float4 of h1 per thread of a group");
  static_assert(kMaxNets == 3, "per-net pointers are read through selects");
  // (a dynamically indexed kernel-parameter array would be copied to local memory)
  auto pick = [&](auto const& a) { return z == 0 ? a[0] : z == 1 ? a[1] : a[2]; };
  const float* w2z = pick(g.w2);

  // ---- (0) operands that were final before the preceding kernel started
  const int act = g.action[row];
  float4 bias1 = make_float4(0.f, 0.f, 0.f, 0.f);
  if (on && g.fc1_part != nullptr) bias1 = __ldg(reinterpret_cast<const float4*>(pick(g.b1)) + c);
  constexpr int kGroup = 6;   // actions whose fc2 weights are fetched together (one L2 round trip per group)
  float wv[kGroup][4];
#pragma unroll
  for (int u = 0; u < kGroup; ++u) {
    const bool ld = on && u < A;
#pragma unroll
    for (int j = 0; j < 4; ++j) wv[u][j] = ld ? __ldg(w2z + (size_t)(4 * c + j) * A + u) : 0.f;
  }
  float b2v = 0.f;
  if (t < nets * A) {
    const int zz = t / A, a = t - zz * A;
    b2v = (zz == 0 ? g.b2[0] : zz == 1 ? g.b2[1] : g.b2[2])[g.shared_bias ? 0 : a];
  }
  float rew = 0.f, disc = 0.f, wrow = 1.0f;
  if (t < 32) {
    rew = g.reward[row]; disc = g.discount[row];
    if (g.row_weight) wrow = g.row_weight[row];
  }
  float w2a[2];   // fc2 weights of the taken action for the dh1 columns this thread writes
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int k = t + i * kHeadThreads;
    w2a[i] = k < kFc1Out ? __ldg(g.w2[0] + (size_t)k * A + act) : 0.f;
  }
  if (t < kMaxNets * kMaxActions) (&q_s[0][0])[t] = -INFINITY;
  asm volatile("griddepcontrol.wait;" ::: "memory");
  // the 32 rows of a small batch leave 116 SMs idle: let fc1's dgrad set up there and pull its weights meanwhile
  asm volatile("griddepcontrol.launch_dependents;");

  // ---- (1) h1: thread c of group z owns columns 4c..4c+3 of net z
  float4 hv = make_float4(0.f, 0.f, 0.f, 0.f);
  if (on) {
    if (g.fc1_part != nullptr) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      const size_t split_stride = (size_t)g.batch * (kFc1Out / 4);
      const float4* src = reinterpret_cast<const float4*>(g.fc1_part + (size_t)row * kFc1Out) + c +
                          (size_t)z * g.fc1_splits * split_stride;
      // all partials of the thread in flight (18 splits at batch 32: one round trip); the additions keep split order
      for (int q0 = 0; q0 < g.fc1_splits; q0 += kDepth) {
        float4 x[kDepth];
#pragma unroll
        for (int u = 0; u < kDepth; ++u)
          if (q0 + u < g.fc1_splits) x[u] = __ldcg(src + (size_t)(q0 + u) * split_stride);
#pragma unroll
        for (int u = 0; u < kDepth; ++u)
          if (q0 + u < g.fc1_splits) { acc.x += x[u].x; acc.y += x[u].y; acc.z += x[u].z; acc.w += x[u].w; }
      }
      hv = make_float4(fmaxf(acc.x + bias1.x, 0.f), fmaxf(acc.y + bias1.y, 0.f), fmaxf(acc.z + bias1.z, 0.f),
                       fmaxf(acc.w + bias1.w, 0.f));
      reinterpret_cast<float4*>(pick(g.h1_out) + (size_t)row * kFc1Out)[c] = hv;
    } else {
      hv = reinterpret_cast<const float4*>(pick(g.h1) + (size_t)row * kFc1Out)[c];
    }
  }
  if (z == 0) reinterpret_cast<float4*>(h_s)[c] = hv;   // relu mask for dh1 below
  // ---- (2) fc2: q[z][a] = sum_k h1[z][k] w2[z][k][a].  Thread c covers k = 4c..4c+3, whose weights are 4A
  // consecutive floats (coalesced across the group); warp shuffle, then the group's 4 warp partials.
  for (int a0 = 0; a0 < A; a0 += kGroup) {
    if (a0 > 0) {   // (more than kGroup actions: the later groups were not prefetched)
#pragma unroll
      for (int u = 0; u < kGroup; ++u) {
        const bool ld = on && a0 + u < A;
#pragma unroll
        for (int j = 0; j < 4; ++j) wv[u][j] = ld ? __ldg(w2z + (size_t)(4 * c + j) * A + a0 + u) : 0.f;
      }
    }
    float part[kGroup];
#pragma unroll
    for (int u = 0; u < kGroup; ++u) {
      float s = hv.x * wv[u][0];
      s = fmaf(hv.y, wv[u][1], s);
      s = fmaf(hv.z, wv[u][2], s);
      s = fmaf(hv.w, wv[u][3], s);
      part[u] = s;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
      for (int u = 0; u < kGroup; ++u) part[u] += __shfl_xor_sync(0xffffffffu, part[u], o);
    }
    if (lane == 0 && on) {
#pragma unroll
      for (int u = 0; u < kGroup; ++u)
        if (a0 + u < A) red_s[z][gw][a0 + u] = part[u];
    }
  }
  __syncthreads();
  if (t < nets * A) {
    const int zz = t / A, a = t - zz * A;
    float s = b2v;
#pragma unroll
    for (int w = 0; w < kHeadGroup / 32; ++w) s += red_s[zz][w][a];
    q_s[zz][a] = s;
    (zz == 0 ? g.q[0] : zz == 1 ? g.q[1] : g.q[2])[(size_t)row * A + a] = s;
  }
  __syncthreads();
  // ---- (3) target, td error, clipped cotangent
  if (t < 32) {
    float best; int sel;
    warp_argmax(q_s[g.double_q ? 2 : 1][lane], lane, best, sel);   // selector net (lanes >= A hold -inf)
    const float q_next = q_s[1][sel];                                // value from the target net
    const float q_taken = q_s[0][act];
    const float td = rew + disc * q_next - q_taken;
    const float win = g.weight_inside ? wrow : 1.0f;
    float cot = fminf(fmaxf(td * g.inv_mean * win, -g.grad_bound), g.grad_bound);
    if (!g.weight_inside) cot *= wrow;
    const float dq = -cot;
    if (lane == 0) {
      g.td[row] = td;
      g.row_loss[row] = 0.5f * td * td * win;
      g.dq_taken[row] = dq;
      dq_s = dq;
    }
  }
  __syncthreads();
  // ---- (4) dh1 = relu'(h1) * dq * w2[:, act]
  const float dq = dq_s;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int k = t + i * kHeadThreads;
    if (k < kFc1Out) {
      const float d = h_s[k] > 0.f ? dq * w2a[i] : 0.f;
      g.dh1[(size_t)row * kFc1Out + k] = g.round_tf32 ? tf32_rn(d) : d;
    }
  }
  tl_end(g.tl, tl_t0);
}
inline auto head_td_for(int batch) { return batch <= 64 ? head_td_kernel<20> : head_td_kernel<4>; }

// fc2 parameter gradients + scalar loss.  thread (k,a): gW2[k,a] = sum_b h1[b,k] * dq[b,a];
// block 0 also writes the bias gradient and loss (fixed summation order => reproducible).
// Host-visible loss ring: every update writes (update sequence number, mean loss) into slot
// seq % 8 of a pinned, device-mapped array, so the host can pick a step's loss up by polling plain
// memory - no copy-engine operation and no stream synchronisation on the training stream.
constexpr int kLossRing = 8;
struct LossRing {
  unsigned long long* ring;   // pinned host memory (device-accessible); null = off
  unsigned* seq;              // device counter of updates executed
};

__global__ void fc2_wgrad_kernel(int batch, int A, int shared_bias, float inv_mean, const float* __restrict__ h1,
                                 const int32_t* __restrict__ action, const float* __restrict__ dq_taken,
                                 const float* __restrict__ row_loss, float* __restrict__ gw2,
                                 float* __restrict__ gb2, float* __restrict__ loss, LossRing ring,
                                 unsigned long long* tl) {
  tl_begin(tl);
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e < kFc1Out * A) {
    int k = e / A, a = e - k * A;
    float s = 0.f;
    for (int b = 0; b < batch; ++b)
      if (action[b] == a) s = fmaf(h1[(size_t)b * kFc1Out + k], dq_taken[b], s);
    gw2[e] = s;
  }
  if (blockIdx.x == 0 && threadIdx.x < 32) {
    int lane = threadIdx.x;
    if (shared_bias) {
      float s = 0.f;
      for (int b = lane; b < batch; b += 32) s += dq_taken[b];
      s = warp_sum(s);
      if (lane == 0) gb2[0] = s;
    } else if (lane < A) {
      float s = 0.f;
      for (int b = 0; b < batch; ++b)
        if (action[b] == lane) s += dq_taken[b];
      gb2[lane] = s;
    }
    float l = 0.f;
    for (int b = lane; b < batch; b += 32) l += row_loss[b];
    l = warp_sum(l);
    if (lane == 0) {
      const float mean = l * inv_mean;
      *loss = mean;
      if (ring.ring != nullptr) {   // (seq, loss) as one 8-byte store into pinned host memory
        const unsigned seq = *ring.seq + 1;
        *ring.seq = seq;
        ring.ring[seq & (kLossRing - 1)] = ((unsigned long long)seq << 32) | __float_as_uint(mean);
        __threadfence_system();
      }
    }
  }
  tl_end(tl);
}

// Large batches: the serial batch loop above becomes the critical path, so slice q of the batch
// (blockIdx.y) leaves its partial in part[q][512*A] and the optimiser kernel sums the slices in order.
// Block (0, 0) also produces the bias gradient and the loss exactly like fc2_wgrad_kernel.
__global__ void fc2_wgrad_sliced_kernel(int batch, int A, int shared_bias, float inv_mean, const float* __restrict__ h1,
                                        const int32_t* __restrict__ action, const float* __restrict__ dq_taken,
                                        const float* __restrict__ row_loss, float* __restrict__ part,
                                        float* __restrict__ gb2, float* __restrict__ loss, LossRing ring,
                                        unsigned long long* tl) {
  tl_begin(tl);
  const int n = kFc1Out * A, e = blockIdx.x * blockDim.x + threadIdx.x;
  const int q = blockIdx.y, nq = gridDim.y;
  const int beg = (int)((int64_t)batch * q / nq), end = (int)((int64_t)batch * (q + 1) / nq);
  if (e < n) {
    int k = e / A, a = e - k * A;
    float s = 0.f;
#pragma unroll 4
    for (int b = beg; b < end; ++b)
      if (__ldg(action + b) == a) s = fmaf(__ldg(h1 + (size_t)b * kFc1Out + k), __ldg(dq_taken + b), s);
    part[(size_t)q * n + e] = s;
  }
  if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x < 32) {
    int lane = threadIdx.x;
    if (shared_bias) {
      float s = 0.f;
      for (int b = lane; b < batch; b += 32) s += dq_taken[b];
      s = warp_sum(s);
      if (lane == 0) gb2[0] = s;
    } else if (lane < A) {
      float s = 0.f;
      for (int b = 0; b < batch; ++b)
        if (action[b] == lane) s += dq_taken[b];
      gb2[lane] = s;
    }
    float l = 0.f;
    for (int b = lane; b < batch; b += 32) l += row_loss[b];
    l = warp_sum(l);
    if (lane == 0) {
      const float mean = l * inv_mean;
      *loss = mean;
      if (ring.ring != nullptr) {
        const unsigned seq = *ring.seq + 1;
        *ring.seq = seq;
        ring.ring[seq & (kLossRing - 1)] = ((unsigned long long)seq << 32) | __float_as_uint(mean);
        __threadfence_system();
      }
    }
  }
  tl_end(tl);
}

// K5: optax.rmsprop(centered=True) + apply_updates, one pass over the flat arena
// (op order of Digraph.gv:711-752; eps inside the rsqrt).
__global__ void rmsprop_kernel(float* __restrict__ p, float* __restrict__ mu, float* __restrict__ nu,
                               const float* __restrict__ g, int64_t n, float lr, float decay, float eps,
                               unsigned long long* tl) {
  tl_begin(tl);
  const float om = 1.0f - decay;
  int64_t i4 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t n4 = n >> 2;
  if (i4 < n4) {
    float4 gv = reinterpret_cast<const float4*>(g)[i4];
    float4 mv = reinterpret_cast<float4*>(mu)[i4];
    float4 nv = reinterpret_cast<float4*>(nu)[i4];
    float4 pv = reinterpret_cast<float4*>(p)[i4];
#define DQNB_RMS(c)                                              \
  mv.c = om * gv.c + decay * mv.c;                               \
  nv.c = om * (gv.c * gv.c) + decay * nv.c;                      \
  pv.c = pv.c + (-lr) * (gv.c * rsqrtf(nv.c - mv.c * mv.c + eps));
    DQNB_RMS(x) DQNB_RMS(y) DQNB_RMS(z) DQNB_RMS(w)
#undef DQNB_RMS
    reinterpret_cast<float4*>(mu)[i4] = mv;
    reinterpret_cast<float4*>(nu)[i4] = nv;
    reinterpret_cast<float4*>(p)[i4] = pv;
  }
  if (i4 == n4) {   // scalar tail
    for (int64_t i = n4 << 2; i < n; ++i) {
      float gv = g[i];
      float m = om * gv + decay * mu[i];
      float v = om * (gv * gv) + decay * nu[i];
      mu[i] = m; nu[i] = v;
      p[i] = p[i] + (-lr) * (gv * rsqrtf(v - m * m + eps));
    }
  }
  tl_end(tl);
}

// K5 with the split-K reduction of the weight gradients folded in.  A region of the arena whose
// gradient still sits in `splits` partial buffers (each `count` floats, arena layout) is summed
// here in split order (bit-reproducible) and written to g; other elements read g as is.
// APPLY=false only materialises g (the data-parallel learner all-reduces it before the update).
struct GradRegion {
  int64_t off, count;      // floats; both multiples of 4
  const float* part;       // split q's partial of this region starts at part + q * stride
  int64_t stride;
  int splits;
};
constexpr int kMaxGradRegions = 8;
struct GradRegions {
  GradRegion r[kMaxGradRegions];
  int n;
};

__device__ __forceinline__ void rmsprop4(float4& pv, float4& mv, float4& nv, const float4& gv, float lr, float decay,
                                         float eps) {
  const float om = 1.0f - decay;
#define DQNB_RMS(c)                                              \
  mv.c = om * gv.c + decay * mv.c;                               \
  nv.c = om * (gv.c * gv.c) + decay * nv.c;                      \
  pv.c = pv.c + (-lr) * (gv.c * rsqrtf(nv.c - mv.c * mv.c + eps));
  DQNB_RMS(x) DQNB_RMS(y) DQNB_RMS(z) DQNB_RMS(w)
#undef DQNB_RMS
}

// Elements [0, split_end) get 8 lanes per float4: lane s sums partials s, s+8, ... and an xor
// butterfly adds the 8 lane sums (every lane performs the same commutative additions, so the
// result is reproducible).  Elements [skip_lo, skip_hi) are updated elsewhere (fc1's fused
// weight-gradient epilogue) and are not visited.  The rest: one thread per float4, g read as is.
// fp16, transposed (K-major) copies of the convolution filters for the fp16-operand forward GEMMs: the arena
// holds a layer's filters as [K][N] fp32 (HWIO), the GEMM wants h16[n][k].  Up to three layers (arena offset,
// K, N, destination); whoever writes the weights keeps the copies current.  See H16ConvFwd / H16ConvFwdDoubleQ.
struct H16Shadow {
  static constexpr int kLayers = 3;
  __half* dst[kLayers];     // null = off
  int64_t off[kLayers];
  int K[kLayers], N[kLayers];
  // one more range kept as a plain (same-layout) fp16 copy: fc1's weights [3136][512], the MN-major B operand of the
  // fp16 fc1 forward and never transposed
  __half* plain_dst;        // null = off
  int64_t plain_off, plain_count;
  __device__ __forceinline__ void put4(int64_t i, const float4& v) const {
    if (plain_dst != nullptr && i >= plain_off && i < plain_off + plain_count) {
      const __half2 lo = __floats2half2_rn(v.x, v.y), hi = __floats2half2_rn(v.z, v.w);
      uint2 u;
      u.x = *reinterpret_cast<const uint32_t*>(&lo);
      u.y = *reinterpret_cast<const uint32_t*>(&hi);
      *reinterpret_cast<uint2*>(plain_dst + (i - plain_off)) = u;
    }
#pragma unroll
    for (int l = 0; l < kLayers; ++l) {
      if (dst[l] == nullptr || i < off[l] || i >= off[l] + (int64_t)K[l] * N[l]) continue;
      const int e = (int)(i - off[l]), k = e / N[l], n = e - k * N[l];
      __half* d = dst[l] + (size_t)n * K[l] + k;
      d[0] = __float2half_rn(v.x);
      d[(size_t)K[l]] = __float2half_rn(v.y);
      d[(size_t)2 * K[l]] = __float2half_rn(v.z);
      d[(size_t)3 * K[l]] = __float2half_rn(v.w);
    }
  }
};
inline H16Shadow h16_none() {
  H16Shadow s;
  for (int l = 0; l < H16Shadow::kLayers; ++l) { s.dst[l] = nullptr; s.off[l] = 0; s.K[l] = 0; s.N[l] = 0; }
  s.plain_dst = nullptr; s.plain_off = s.plain_count = 0;
  return s;
}
template <bool APPLY>
__global__ void rmsprop_reduce_kernel(float* __restrict__ p, float* __restrict__ p_r, H16Shadow h16, float* __restrict__ mu,
                                      float* __restrict__ nu, float* __restrict__ g, int64_t n, GradRegions regs, int64_t split_end,
                                      int64_t skip_lo, int64_t skip_hi, float lr, float decay, float eps,
                                      unsigned long long* tl) {
  tl_begin(tl);
  // (launched as a programmatic dependent of the last weight-gradient kernel of the chain: placed while that one
  // still stores, released when it has completed; a no-op for ordinary launches)
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nA = (split_end >> 2) * 8;
  if (gid < nA) {
    const int64_t i4 = gid >> 3, i = i4 << 2;
    const int sub = (int)(gid & 7);
    float4 gv = make_float4(0.f, 0.f, 0.f, 0.f);
    bool summed = false;
#pragma unroll
    for (int r = 0; r < kMaxGradRegions; ++r) {
      if (r < regs.n && i >= regs.r[r].off && i < regs.r[r].off + regs.r[r].count) {
        const float* src = regs.r[r].part + (i - regs.r[r].off);
        // four partials in flight per round trip (conv1's weight gradient leaves ~100 of them: 13 per lane, which
        // a rolled loop fetched one L2 latency after the other); the additions keep their q order
        const int ns = regs.r[r].splits;
        const size_t st = regs.r[r].stride;
        int q = sub;
        for (; q + 24 < ns; q += 32) {
          const float4 t0 = __ldcg(reinterpret_cast<const float4*>(src + (size_t)q * st));
          const float4 t1 = __ldcg(reinterpret_cast<const float4*>(src + (size_t)(q + 8) * st));
          const float4 t2 = __ldcg(reinterpret_cast<const float4*>(src + (size_t)(q + 16) * st));
          const float4 t3 = __ldcg(reinterpret_cast<const float4*>(src + (size_t)(q + 24) * st));
          gv.x += t0.x; gv.y += t0.y; gv.z += t0.z; gv.w += t0.w;
          gv.x += t1.x; gv.y += t1.y; gv.z += t1.z; gv.w += t1.w;
          gv.x += t2.x; gv.y += t2.y; gv.z += t2.z; gv.w += t2.w;
          gv.x += t3.x; gv.y += t3.y; gv.z += t3.z; gv.w += t3.w;
        }
        for (; q < ns; q += 8) {
          const float4 t = __ldcg(reinterpret_cast<const float4*>(src + (size_t)q * st));
          gv.x += t.x; gv.y += t.y; gv.z += t.z; gv.w += t.w;
        }
        summed = true;
      }
    }
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) {
      gv.x += __shfl_xor_sync(0xffffffffu, gv.x, o);
      gv.y += __shfl_xor_sync(0xffffffffu, gv.y, o);
      gv.z += __shfl_xor_sync(0xffffffffu, gv.z, o);
      gv.w += __shfl_xor_sync(0xffffffffu, gv.w, o);
    }
    if (sub == 0) {
      if (summed) reinterpret_cast<float4*>(g)[i4] = gv;
      else gv = reinterpret_cast<const float4*>(g)[i4];
      if (APPLY) {
        float4 mv = reinterpret_cast<float4*>(mu)[i4], nv = reinterpret_cast<float4*>(nu)[i4];
        float4 pv = reinterpret_cast<float4*>(p)[i4];
        rmsprop4(pv, mv, nv, gv, lr, decay, eps);
        reinterpret_cast<float4*>(mu)[i4] = mv;
        reinterpret_cast<float4*>(nu)[i4] = nv;
        reinterpret_cast<float4*>(p)[i4] = pv;
        if (p_r != nullptr) reinterpret_cast<float4*>(p_r)[i4] = tf32_rn4(pv);
        h16.put4(i, pv);
      }
    }
  } else if (APPLY) {
    int64_t i = split_end + ((gid - nA) << 2);
    if (i >= skip_lo) i += skip_hi - skip_lo;
    const int64_t n4 = n >> 2, i4 = i >> 2;
    if (i4 < n4) {
      const float4 gv = reinterpret_cast<const float4*>(g)[i4];
      float4 mv = reinterpret_cast<float4*>(mu)[i4], nv = reinterpret_cast<float4*>(nu)[i4];
      float4 pv = reinterpret_cast<float4*>(p)[i4];
      rmsprop4(pv, mv, nv, gv, lr, decay, eps);
      reinterpret_cast<float4*>(mu)[i4] = mv;
      reinterpret_cast<float4*>(nu)[i4] = nv;
      reinterpret_cast<float4*>(p)[i4] = pv;
      if (p_r != nullptr) reinterpret_cast<float4*>(p_r)[i4] = tf32_rn4(pv);
      h16.put4(i, pv);
    } else if (i4 == n4) {   // scalar tail
      const float om = 1.0f - decay;
      for (int64_t k = n4 << 2; k < n; ++k) {
        float gk = g[k];
        float m = om * gk + decay * mu[k];
        float v = om * (gk * gk) + decay * nu[k];
        mu[k] = m; nu[k] = v;
        p[k] = p[k] + (-lr) * (gk * rsqrtf(v - m * m + eps));
        if (p_r != nullptr) p_r[k] = tf32_rn(p[k]);
      }
    }
  }
  tl_end(tl);
}
// Sums the split-K partials of the arena range [lo, hi) (multiples of 4) into g, with the very additions of
// rmsprop_reduce_kernel (8 lanes per float4, xor butterfly): the data-parallel learner materialises the gradient
// in two pieces so that the first (fc1 + fc2, 95 % of the bytes) can be all-reduced while the convolution
// backward still runs.  Elements without a registered region keep what the producing kernel wrote into g.
__global__ void grad_reduce_range_kernel(float* __restrict__ g, GradRegions regs, int64_t lo, int64_t hi,
                                         unsigned long long* tl) {
  tl_begin(tl);
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t i4 = (lo >> 2) + (gid >> 3), i = i4 << 2;
  const int sub = (int)(gid & 7);
  const bool live = i < hi;
  float4 gv = make_float4(0.f, 0.f, 0.f, 0.f);
  bool summed = false;
  if (live) {
#pragma unroll
    for (int r = 0; r < kMaxGradRegions; ++r) {
      if (r < regs.n && i >= regs.r[r].off && i < regs.r[r].off + regs.r[r].count) {
        const float* src = regs.r[r].part + (i - regs.r[r].off);
        // four partials in flight per round trip (conv1's weight gradient leaves ~100 of them: 13 per lane, which
        // a rolled loop fetched one L2 latency after the other); the additions keep their q order
        const int ns = regs.r[r].splits;
        const size_t st = regs.r[r].stride;
        int q = sub;
        for (; q + 24 < ns; q += 32) {
          const float4 t0 = __ldcg(reinterpret_cast<const float4*>(src + (size_t)q * st));
          const float4 t1 = __ldcg(reinterpret_cast<const float4*>(src + (size_t)(q + 8) * st));
          const float4 t2 = __ldcg(reinterpret_cast<const float4*>(src + (size_t)(q + 16) * st));
          const float4 t3 = __ldcg(reinterpret_cast<const float4*>(src + (size_t)(q + 24) * st));
          gv.x += t0.x; gv.y += t0.y; gv.z += t0.z; gv.w += t0.w;
          gv.x += t1.x; gv.y += t1.y; gv.z += t1.z; gv.w += t1.w;
          gv.x += t2.x; gv.y += t2.y; gv.z += t2.z; gv.w += t2.w;
          gv.x += t3.x; gv.y += t3.y; gv.z += t3.z; gv.w += t3.w;
        }
        for (; q < ns; q += 8) {
          const float4 t = __ldcg(reinterpret_cast<const float4*>(src + (size_t)q * st));
          gv.x += t.x; gv.y += t.y; gv.z += t.z; gv.w += t.w;
        }
        summed = true;
      }
    }
  }
#pragma unroll
  for (int o = 4; o > 0; o >>= 1) {
    gv.x += __shfl_xor_sync(0xffffffffu, gv.x, o);
    gv.y += __shfl_xor_sync(0xffffffffu, gv.y, o);
    gv.z += __shfl_xor_sync(0xffffffffu, gv.z, o);
    gv.w += __shfl_xor_sync(0xffffffffu, gv.w, o);
  }
  if (live && summed && sub == 0) reinterpret_cast<float4*>(g)[i4] = gv;
  tl_end(tl);
}
inline unsigned grad_reduce_range_blocks(int64_t lo, int64_t hi) {
  return (unsigned)((((hi - lo + 3) >> 2) * 8 + 255) / 256);
}
__global__ void round_tf32_kernel(const float* __restrict__ src, float* __restrict__ dst, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = tf32_rn(src[i]);
}
inline unsigned rmsprop_reduce_blocks(int64_t n, int64_t split_end, int64_t skip_lo, int64_t skip_hi) {
  int64_t threads = (split_end >> 2) * 8 + ((n - split_end - (skip_hi - skip_lo)) >> 2) + 1;
  return (unsigned)((threads + 255) / 256);
}

// MGSC adjoint dots (dqn_mgsc_batched/agent.py:171-214): for row i of the last update,
//   out[i] = <v, d loss_i / d theta> = sum over layers l of <dy_l[i], V_l * x_l[i] + vb_l>
// with dy_l the saved pre-activation cotangents of the backward pass and z_l = V_l * x_l + vb_l the
// layer-local forward products of the direction v (z1..z4, computed by the forward GEMMs with v in
// the weight slot and no ReLU).  fc2's term is formed here.  One CTA per row, float64 partials,
// fixed summation order.
struct AdjointArgs {
  int num_actions, shared_bias;
  const float *dy1, *z1, *dy2, *z2, *dy3, *z3, *dh1, *z4;   // [batch, n_l]
  int n1, n2, n3, n4;
  const float* h1;          // [batch,512] relu(fc1) of the online net
  const float* dq_taken;    // [batch]
  const int32_t* action;
  const float* v_w2;        // [512,A] slice of v
  const float* v_b2;        // [A] or [1]
  float* out;               // [batch]
};

__device__ __forceinline__ double dot_span(const float* __restrict__ a, const float* __restrict__ b, int n, int t) {
  double acc = 0.0;
  const float4* a4 = reinterpret_cast<const float4*>(a);
  const float4* b4 = reinterpret_cast<const float4*>(b);
  for (int i = t; i < (n >> 2); i += 256) {
    const float4 x = a4[i], y = b4[i];
    acc += (double)x.x * y.x + (double)x.y * y.y + (double)x.z * y.z + (double)x.w * y.w;
  }
  return acc;
}

__global__ void __launch_bounds__(256) adjoint_dot_kernel(AdjointArgs g) {
  __shared__ double red[256];
  const int row = blockIdx.x, t = threadIdx.x;
  double acc = dot_span(g.dy1 + (size_t)row * g.n1, g.z1 + (size_t)row * g.n1, g.n1, t);
  acc += dot_span(g.dy2 + (size_t)row * g.n2, g.z2 + (size_t)row * g.n2, g.n2, t);
  acc += dot_span(g.dy3 + (size_t)row * g.n3, g.z3 + (size_t)row * g.n3, g.n3, t);
  acc += dot_span(g.dh1 + (size_t)row * g.n4, g.z4 + (size_t)row * g.n4, g.n4, t);
  const int act = g.action[row];
  const double dq = (double)g.dq_taken[row];
  for (int k = t; k < g.n4; k += 256)
    acc += dq * (double)g.h1[(size_t)row * g.n4 + k] * (double)g.v_w2[(size_t)k * g.num_actions + act];
  if (t == 0) acc += dq * (double)g.v_b2[g.shared_bias ? 0 : act];
  red[t] = acc;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) { if (t < s) red[t] += red[t + s]; __syncthreads(); }
  if (t == 0) g.out[row] = (float)red[0];
}

}  // namespace dqnb
