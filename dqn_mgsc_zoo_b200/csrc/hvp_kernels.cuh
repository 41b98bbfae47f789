// Small kernels of the Hessian-vector product (dqnb_learner_hvp): the second-order MGSC objective of
// dqn_mgsc_batched_reservoir/agent.py:160-198 differentiates THROUGH jax.grad(loss_fn)(expected_online_params, ...),
// i.e. needs H . w with H the Jacobian of the gradient.  Forward-over-reverse (R-op): ReLU masks are locally constant,
//   tangent forward :  zdot_l = W_l * xdot_{l-1} + Wdot_l * x_{l-1} + bdot_l ,  xdot_l = mask_l . zdot_l
//   tangent backward:  ddot_{l-1} = mask . (W_l^T ddot_l + Wdot_l^T d_l)
//                      (grad W_l)dot = x_{l-1}^T ddot_l + xdot_{l-1}^T d_l ,  (grad b_l)dot = sum ddot_l
// Every product above is one of the existing layer GEMMs (gemm_simt<P>) with the direction in a weight or activation
// slot; the kernels here are the glue: sums with masks, the fc2 / TD head, and the final assembly.
#pragma once
#include "head_kernels.cuh"

namespace dqnb {

// out = (act > 0) ? (u + v) : 0      (v may be null; act null = no mask)
__global__ void hvp_add_mask_kernel(const float* __restrict__ u, const float* __restrict__ v,
                                    const float* __restrict__ act, float* __restrict__ out, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float s = u[i] + (v ? v[i] : 0.f);
  out[i] = (act == nullptr || act[i] > 0.f) ? s : 0.f;
}

struct HvpHeadArgs {
  int num_actions, shared_bias, weight_inside;
  float grad_bound, inv_mean;
  const float* h1;        // [batch,512] relu(fc1) of the online net at the last update
  const float* h1dot;     // [batch,512] its tangent (already masked)
  const float* w2;        // [512,A] fc2 weights        b2 is not needed (constant in q, no tangent)
  const float* v_w2;      // [512,A] direction, fc2 weights
  const float* v_b2;      // [A] or [1] direction, fc2 bias
  const int32_t* action;
  const float* td;        // [batch] td errors of the last update
  const float* dq_taken;  // [batch] d loss / d q[b, a_b] of the last update
  const float* row_weight;   // may be null
  float* dq_dot;          // [batch] tangent of dq_taken
  float* dh1_dot;         // [batch,512] tangent of the fc1 pre-activation cotangent (masked)
};

// One CTA per batch row: qdot[a_b], the tangent of the clipped cotangent, and the fc1 cotangent tangent.
//   cot = clip(td * inv_mean * w_in, +-bound) [* w_out];  d cot = 1[|.| < bound] * (-qdot[a]) * inv_mean * w_in [* w_out]
//   (rlax.clip_gradient's backward clip(g) differentiated: 1 inside, 0 outside.)
__global__ void __launch_bounds__(128) hvp_head_kernel(HvpHeadArgs g) {
  __shared__ float red[128];
  __shared__ float dqd_s;
  const int row = blockIdx.x, t = threadIdx.x, A = g.num_actions;
  const int act = g.action[row];
  float acc = 0.f;
  for (int k = t; k < kFc1Out; k += 128) {
    const size_t i = (size_t)row * kFc1Out + k;
    acc = fmaf(g.w2[(size_t)k * A + act], g.h1dot[i], acc);
    acc = fmaf(g.v_w2[(size_t)k * A + act], g.h1[i], acc);
  }
  red[t] = acc;
  __syncthreads();
  for (int s = 64; s > 0; s >>= 1) { if (t < s) red[t] += red[t + s]; __syncthreads(); }
  if (t == 0) {
    const float qdot = red[0] + g.v_b2[g.shared_bias ? 0 : act];
    const float wrow = g.row_weight ? g.row_weight[row] : 1.0f;
    const float win = g.weight_inside ? wrow : 1.0f;
    const float pre = g.td[row] * g.inv_mean * win;
    float cdot = (fabsf(pre) < g.grad_bound) ? (-qdot) * g.inv_mean * win : 0.f;
    if (!g.weight_inside) cdot *= wrow;
    dqd_s = -cdot;
    g.dq_dot[row] = -cdot;
  }
  __syncthreads();
  const float dqd = dqd_s, dq = g.dq_taken[row];
  for (int k = t; k < kFc1Out; k += 128) {
    const size_t i = (size_t)row * kFc1Out + k;
    const float d = g.w2[(size_t)k * A + act] * dqd + g.v_w2[(size_t)k * A + act] * dq;
    g.dh1_dot[i] = g.h1[i] > 0.f ? d : 0.f;
  }
}

// (grad fc2.w)dot[k,a] = sum_{b: a_b = a} (h1dot[b,k] dq_b + h1[b,k] dqdot_b);  (grad fc2.b)dot = sum dqdot
__global__ void hvp_fc2_kernel(int batch, int A, int shared_bias, const float* __restrict__ h1,
                               const float* __restrict__ h1dot, const int32_t* __restrict__ action,
                               const float* __restrict__ dq, const float* __restrict__ dqd, float* __restrict__ gw2,
                               float* __restrict__ gb2) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e < kFc1Out * A) {
    const int k = e / A, a = e - k * A;
    float s = 0.f;
    for (int b = 0; b < batch; ++b)
      if (action[b] == a) {
        s = fmaf(h1dot[(size_t)b * kFc1Out + k], dq[b], s);
        s = fmaf(h1[(size_t)b * kFc1Out + k], dqd[b], s);
      }
    gw2[e] = s;
  }
  if (blockIdx.x == 0 && threadIdx.x < (shared_bias ? 1 : A)) {
    const int a = threadIdx.x;
    float s = 0.f;
    for (int b = 0; b < batch; ++b)
      if (shared_bias || action[b] == a) s += dqd[b];
    gb2[a] = s;
  }
}

// out[i] = a[i] + (i in a bias range ? 0 : b[i]) over [0, n): `a` holds the wgrads with the original inputs and the
// tangent cotangents (weights AND biases), `b` the wgrads with the tangent inputs and the original cotangents, whose
// ones-row "bias gradient" is not part of the product.  Bias ranges are [lo[r], hi[r]).
struct HvpBiasRanges { int64_t lo[4], hi[4]; };
__global__ void hvp_combine_kernel(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ out,
                                   int64_t n, HvpBiasRanges r) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  bool bias = false;
#pragma unroll
  for (int q = 0; q < 4; ++q) bias |= i >= r.lo[q] && i < r.hi[q];
  out[i] = a[i] + (bias ? 0.f : b[i]);
}

}  // namespace dqnb
