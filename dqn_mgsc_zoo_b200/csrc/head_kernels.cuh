// K4 (TD / loss, warp-shuffle reductions), fc2 forward+backward, K5 fused centered RMSProp.
#pragma once
#include "nature_problems.cuh"

namespace dqnb {

constexpr int kMaxActions = 32;   // one lane per action in the warp-level argmax

struct HeadArgs {
  int batch, num_actions, shared_bias, double_q;
  float grad_bound;        // rlax.clip_gradient bound (dqn/agent.py:101-103)
  float inv_mean;          // 1 / (number of rows the loss mean runs over)
  const float* h1[kMaxNets];   // [batch,512] relu(fc1): online(s_tm1), target(s_t), online(s_t)
  const float* w2[kMaxNets];   // [512,A]
  const float* b2[kMaxNets];   // [A] or [1]
  const int32_t* action;
  const float* reward;
  const float* discount;
  const float* row_weight;  // optional per-row weight on the cotangent (MGSC), may be null
  float* q[kMaxNets];       // [batch,A] outputs
  float* td;                // [batch]
  float* row_loss;          // [batch] 0.5*td^2
  float* dq_taken;          // [batch] d loss / d q_tm1[b, a_b]
  float* dh1;               // [batch,512] gradient wrt fc1 pre-activation (relu-masked)
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

// One warp per batch row: fc2 forward for every net, q_learning / double_q_learning target
// (rlax), td, clipped cotangent, per-row loss, and the fc2 input gradient dh1.
__global__ void __launch_bounds__(128) head_td_kernel(HeadArgs g) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= g.batch) return;
  const int A = g.num_actions, nets = g.double_q ? 3 : 2;
  float qv[kMaxNets];   // lane a keeps q_z[row, a]
  for (int z = 0; z < nets; ++z) {
    const float* h = g.h1[z] + (size_t)row * kFc1Out;
    float mine = -INFINITY;
    for (int a = 0; a < A; ++a) {
      float s = 0.f;
      for (int k = lane; k < kFc1Out; k += 32) s = fmaf(h[k], g.w2[z][(size_t)k * A + a], s);
      s = warp_sum(s) + g.b2[z][g.shared_bias ? 0 : a];
      if (lane == a) mine = s;
    }
    qv[z] = mine;
    if (lane < A) g.q[z][(size_t)row * A + lane] = mine;
  }
  float best; int sel;
  warp_argmax(qv[g.double_q ? 2 : 1], lane, best, sel);          // selector net
  const float q_next = __shfl_sync(0xffffffffu, qv[1], sel);      // value from the target net
  const int act = g.action[row];
  const float q_taken = __shfl_sync(0xffffffffu, qv[0], act);
  const float td = g.reward[row] + g.discount[row] * q_next - q_taken;
  float cot = fminf(fmaxf(td * g.inv_mean, -g.grad_bound), g.grad_bound);
  if (g.row_weight) cot *= g.row_weight[row];
  const float dq = -cot;
  if (lane == 0) {
    g.td[row] = td;
    g.row_loss[row] = 0.5f * td * td;
    g.dq_taken[row] = dq;
  }
  const float* h = g.h1[0] + (size_t)row * kFc1Out;
  for (int k = lane; k < kFc1Out; k += 32)
    g.dh1[(size_t)row * kFc1Out + k] = h[k] > 0.f ? dq * g.w2[0][(size_t)k * A + act] : 0.f;
}

// fc2 parameter gradients + scalar loss.  thread (k,a): gW2[k,a] = sum_b h1[b,k] * dq[b,a];
// block 0 also writes the bias gradient and loss (fixed summation order => reproducible).
__global__ void fc2_wgrad_kernel(int batch, int A, int shared_bias, float inv_mean, const float* __restrict__ h1,
                                 const int32_t* __restrict__ action, const float* __restrict__ dq_taken,
                                 const float* __restrict__ row_loss, float* __restrict__ gw2,
                                 float* __restrict__ gb2, float* __restrict__ loss) {
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
    if (lane == 0) *loss = l * inv_mean;
  }
}

// K5: optax.rmsprop(centered=True) + apply_updates, one pass over the flat arena
// (op order of Digraph.gv:711-752; eps inside the rsqrt).
__global__ void rmsprop_kernel(float* __restrict__ p, float* __restrict__ mu, float* __restrict__ nu,
                               const float* __restrict__ g, int64_t n, float lr, float decay, float eps) {
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
}

}  // namespace dqnb
