// K8b: MGSC meta_update (dqn_mgsc_batched/agent.py:152-220) on the device.
//
//   p = softmax(logits)                                        replay_circular.py:69-86
//   g_i = grad of the batch-of-1 loss of meta transition i     agent.py:152-158 (clip acts on td itself)
//   G = sum_i p_i g_i ; theta_exp = theta + u(G; mu, nu)       agent.py:171-179 (virtual RMSProp step)
//   g2 = grad at theta_exp (target = theta) on the online transition ; theta_tgt = theta_exp + u(g2; mu1, nu1)
//   L = |theta_exp - theta_tgt|^2 with theta_tgt a constant    agent.py:183-195 (stop_gradient)
//   dL/dp_i = <v, g_i>,  v = 2 (theta_exp - theta_tgt) * du/dG  (SURVEY R12 identity)
//   dL/dlogit_j = p_j (dL/dp_j - sum_i p_i dL/dp_i) ; Adam on the (M,) logits   agent.py:206-214
//
// The vmapped per-sample gradients are never materialised (the reference holds M x 6.75 MB): an
// internal batch-of-M learner (`bat`: defer_apply, row_weighted, loss over 1 row) runs ONE batched
// forward/backward whose row cotangents are p_i * clip(td_i), so its gradient arena is G; the
// inner products <v, g_i> come from the same pass's saved activations and cotangents
// (dqnb_learner_adjoint_dots: <v, p_i g_i> = sum_l <dy_l[i], V_l x_l[i]>), divided by p_i.
// A batch-of-1 learner (`sub`) takes the gradient at theta_exp on the online transition.
#include <algorithm>
#include <vector>

#include "learner_internal.cuh"

using namespace dqnb;

namespace {
constexpr int kLossBlocks = 128;   // partial sums of the meta loss
constexpr int kMaxMeta = 1024;

__global__ void softmax_kernel(const float* __restrict__ logits, float* __restrict__ p, int M) {
  __shared__ float red[256];
  const int t = threadIdx.x;
  float mx = -INFINITY;
  for (int i = t; i < M; i += 256) mx = fmaxf(mx, logits[i]);
  red[t] = mx;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) { if (t < s) red[t] = fmaxf(red[t], red[t + s]); __syncthreads(); }
  mx = red[0];
  __syncthreads();
  float sum = 0.f;
  for (int i = t; i < M; i += 256) sum += expf(logits[i] - mx);
  red[t] = sum;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) { if (t < s) red[t] += red[t + s]; __syncthreads(); }
  const float lse = mx + logf(red[0]);
  for (int i = t; i < M; i += 256) p[i] = expf(logits[i] - lse);
}

// optax 0.1.2 rmsprop(centered) as a pure function: (theta, G, mu, nu) -> (theta_exp, mu1, nu1).
__global__ void virtual_step_kernel(const float* __restrict__ theta, const float* __restrict__ G,
                                    const float* __restrict__ mu, const float* __restrict__ nu,
                                    float* __restrict__ theta_exp, float* __restrict__ mu1, float* __restrict__ nu1,
                                    int64_t P, float lr, float decay, float eps) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= P) return;
  const float g = G[j], om = 1.0f - decay;
  const float m = om * g + decay * mu[j];
  const float v = om * (g * g) + decay * nu[j];
  mu1[j] = m; nu1[j] = v;
  theta_exp[j] = theta[j] + (-lr) * (g * rsqrtf(v - m * m + eps));
}

// Second virtual step and the meta-loss cotangent:
//   u2 = -lr g2 rsqrt(nu2 - mu2^2 + eps) = theta_tgt - theta_exp ; L = sum u2^2 ; e = -2 u2
//   v = e * du/dG,  du/dG = -lr s (1 - (1-d) G (G - mu1) s^2),  s = rsqrt(nu1 - mu1^2 + eps)
__global__ void meta_v_kernel(const float* __restrict__ g2, const float* __restrict__ G, const float* __restrict__ mu1,
                              const float* __restrict__ nu1, float* __restrict__ v, float* __restrict__ loss_part,
                              int64_t P, float lr, float decay, float eps) {
  __shared__ float red[256];
  const float om = 1.0f - decay;
  float acc = 0.f;
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < P; j += (int64_t)gridDim.x * blockDim.x) {
    const float g = g2[j], m1 = mu1[j], n1 = nu1[j], gg = G[j];
    const float m2 = om * g + decay * m1;
    const float n2 = om * (g * g) + decay * n1;
    const float u2 = (-lr) * (g * rsqrtf(n2 - m2 * m2 + eps));
    acc = fmaf(u2, u2, acc);
    const float s = rsqrtf(n1 - m1 * m1 + eps);
    const float du = (-lr) * s * (1.0f - om * gg * (gg - m1) * s * s);
    v[j] = (-2.0f * u2) * du;
  }
  red[threadIdx.x] = acc;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) { if (threadIdx.x < s) red[threadIdx.x] += red[threadIdx.x + s]; __syncthreads(); }
  if (threadIdx.x == 0) loss_part[blockIdx.x] = red[0];
}

// Second-order objective (dqn_mgsc_batched_reservoir/agent.py:160-198: no stop_gradient on target_online_params):
//   L = |theta_exp - theta_tgt|^2 = sum u2^2,  u2 = -lr g2 s2,  s2 = rsqrt(nu2 - mu2^2 + eps),
//   mu2 = (1-d) g2 + d mu1,  nu2 = (1-d) g2^2 + d nu1,  g2 = grad loss(theta_exp)  - all functions of G.
// This kernel forms the cotangent on g2 (wdir = 2 u2 * du2/dg2, the direction of the Hessian-vector product) and the
// part of dL/dG that flows through mu1 and nu1 (extra); meta_second_finish_kernel adds the part through theta_exp.
__global__ void meta_second_kernel(const float* __restrict__ g2, const float* __restrict__ G, const float* __restrict__ mu1,
                                   const float* __restrict__ nu1, float* __restrict__ wdir, float* __restrict__ extra,
                                   float* __restrict__ loss_part, int64_t P, float lr, float decay, float eps) {
  __shared__ float red[256];
  const float om = 1.0f - decay;
  float acc = 0.f;
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < P; j += (int64_t)gridDim.x * blockDim.x) {
    const float g = g2[j], m1 = mu1[j], n1 = nu1[j], gg = G[j];
    const float m2 = om * g + decay * m1;
    const float n2 = om * (g * g) + decay * n1;
    const float s2 = rsqrtf(n2 - m2 * m2 + eps), s23 = s2 * s2 * s2;
    const float u2 = (-lr) * (g * s2);
    acc = fmaf(u2, u2, acc);
    const float e = 2.0f * u2;
    const float du_dg = (-lr) * (s2 - om * g * (g - m2) * s23);
    const float du_dmu1 = (-lr) * g * decay * m2 * s23;
    const float du_dnu1 = 0.5f * lr * g * decay * s23;
    wdir[j] = e * du_dg;
    extra[j] = e * du_dmu1 * om + e * du_dnu1 * 2.0f * om * gg;
  }
  red[threadIdx.x] = acc;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) { if (threadIdx.x < s) red[threadIdx.x] += red[threadIdx.x + s]; __syncthreads(); }
  if (threadIdx.x == 0) loss_part[blockIdx.x] = red[0];
}
// v = (H wdir) * d theta_exp / dG + extra
__global__ void meta_second_finish_kernel(const float* __restrict__ t, const float* __restrict__ G,
                                          const float* __restrict__ mu1, const float* __restrict__ nu1,
                                          const float* __restrict__ extra, float* __restrict__ v, int64_t P, float lr,
                                          float decay, float eps) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= P) return;
  const float om = 1.0f - decay, m1 = mu1[j], gg = G[j];
  const float s = rsqrtf(nu1[j] - m1 * m1 + eps);
  const float du = (-lr) * s * (1.0f - om * gg * (gg - m1) * s * s);
  v[j] = t[j] * du + extra[j];
}

// softmax backward + optax.adam (scale_by_adam bias-corrected, eps outside the sqrt) on the logits.
// One block; M <= kMaxMeta.
// `wdot[i]` = <v, p_i g_i>: dL/dp_i = wdot[i] / p_i.  A p_i that underflowed leaves dL/dp_i = 0: its
// logit gradient p_i (dL/dp_i - mean) vanishes for any finite dL/dp_i.
__global__ void logits_adam_kernel(const float* __restrict__ logits, const float* __restrict__ p,
                                   const float* __restrict__ wdot, const float* __restrict__ loss_part,
                                   float* __restrict__ dldp, float* __restrict__ adam_m, float* __restrict__ adam_v,
                                   float* __restrict__ out_logits, float* __restrict__ out_loss, int M, int count,
                                   float lr, float b1, float b2, float eps) {
  __shared__ float red[256];
  __shared__ float d_s[kMaxMeta];
  const int t = threadIdx.x;
  float acc = 0.f;
  for (int i = t; i < M; i += 256) {
    const float d = p[i] > 1e-30f ? wdot[i] / p[i] : 0.f;
    d_s[i] = d;
    dldp[i] = d;
    acc = fmaf(p[i], d, acc);
  }
  red[t] = acc;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) { if (t < s) red[t] += red[t + s]; __syncthreads(); }
  const float mean = red[0];
  const float c1 = 1.0f - powf(b1, (float)count), c2 = 1.0f - powf(b2, (float)count);
  for (int i = t; i < M; i += 256) {
    const float g = p[i] * (d_s[i] - mean);
    const float m = (1.0f - b1) * g + b1 * adam_m[i];
    const float v = (1.0f - b2) * (g * g) + b2 * adam_v[i];
    adam_m[i] = m; adam_v[i] = v;
    out_logits[i] = logits[i] + (-lr) * ((m / c1) / (sqrtf(v / c2) + eps));
  }
  if (t == 0) {
    float l = 0.f;
    for (int b = 0; b < kLossBlocks; ++b) l += loss_part[b];
    *out_loss = l;
  }
}
}  // namespace

struct dqnb_mgsc {
  dqnb_learner* main = nullptr;
  dqnb_learner* bat = nullptr;     // batch-of-M, defer_apply + row_weighted: G and the adjoint dots in one pass
  dqnb_learner* sub = nullptr;     // batch-of-1, defer_apply: gradient at theta_exp on the online transition
  dqnb_learner_config lcfg{};
  dqnb_mgsc_config cfg{};
  int device = 0, M = 0;
  int64_t P = 0;
  size_t img = 0;
  // meta batch + online transition on the device
  uint8_t *mb_s0 = nullptr, *mb_s1 = nullptr;
  int32_t* mb_a = nullptr;
  float *mb_r = nullptr, *mb_d = nullptr;
  float *G = nullptr, *theta_exp = nullptr, *mu1 = nullptr, *nu1 = nullptr, *v = nullptr;
  float *wdir = nullptr, *extra = nullptr, *hv = nullptr;      // second-order objective only
  float *logits = nullptr, *p = nullptr, *dldp = nullptr, *dot_part = nullptr, *loss_part = nullptr;
  float *adam_m = nullptr, *adam_v = nullptr, *out = nullptr;   // out: [M] new logits + [1] loss
  float* pinned = nullptr;                                      // [2M + 8]
  uint8_t* pinned_on = nullptr;                                 // [2 * img]: the online transition's stacks
  int32_t count = 0;                                            // Adam step count
};

namespace {
template <class T>
int dalloc(T** p, size_t count) {
  DQNB_CUDA(cudaMalloc(p, sizeof(T) * std::max<size_t>(count, 4)));
  DQNB_CUDA(cudaMemset(*p, 0, sizeof(T) * std::max<size_t>(count, 4)));
  return DQNB_OK;
}

// Steps 3..11 of the header comment; the meta batch (M transitions) and the online transition
// (index M) are already in h->mb_*; h->pinned[0..M) holds the incoming logits.
int run_meta_update(dqnb_mgsc* h, float* new_logits_host, float* meta_loss_host, cudaStream_t st) {
  const int M = h->M;
  const int64_t P = h->P;
  const uintptr_t s = reinterpret_cast<uintptr_t>(st);
  void *theta = nullptr, *theta_t = nullptr, *mu = nullptr, *nu = nullptr, *sub_grad = nullptr, *bat_grad = nullptr;
  int64_t n = 0;
  DQNB_TRY(dqnb_learner_device_ptr(h->main, DQNB_ONLINE, &theta, &n));
  DQNB_TRY(dqnb_learner_device_ptr(h->main, DQNB_TARGET, &theta_t, &n));
  DQNB_TRY(dqnb_learner_device_ptr(h->main, DQNB_MU, &mu, &n));
  DQNB_TRY(dqnb_learner_device_ptr(h->main, DQNB_NU, &nu, &n));
  DQNB_TRY(dqnb_learner_device_ptr(h->sub, DQNB_GRAD, &sub_grad, &n));
  DQNB_TRY(dqnb_learner_device_ptr(h->bat, DQNB_GRAD, &bat_grad, &n));
  DQNB_CUDA(cudaMemcpyAsync(h->logits, h->pinned, sizeof(float) * M, cudaMemcpyHostToDevice, st));
  softmax_kernel<<<1, 256, 0, st>>>(h->logits, h->p, M);
  // G = sum_i p_i g_i at (theta, theta_target): one batched pass with row weights p
  DQNB_TRY(dqnb_learner_write_dev(h->bat, DQNB_ROW_WEIGHT, h->p, M, s));
  DQNB_TRY(dqnb_learner_write_dev(h->bat, DQNB_ONLINE, static_cast<const float*>(theta), P, s));
  DQNB_TRY(dqnb_learner_write_dev(h->bat, DQNB_TARGET, static_cast<const float*>(theta_t), P, s));
  DQNB_TRY(dqnb_learner_update(h->bat, h->mb_s0, h->mb_a, h->mb_r, h->mb_d, h->mb_s1, s));
  DQNB_CUDA(cudaMemcpyAsync(h->G, bat_grad, sizeof(float) * P, cudaMemcpyDeviceToDevice, st));
  const unsigned pb = (unsigned)((P + 255) / 256);
  virtual_step_kernel<<<pb, 256, 0, st>>>(static_cast<const float*>(theta), h->G, static_cast<const float*>(mu),
                                          static_cast<const float*>(nu), h->theta_exp, h->mu1, h->nu1, P,
                                          h->lcfg.learning_rate, h->lcfg.decay, h->lcfg.eps);
  // gradient at theta_exp against the *online* parameters as target, on the online transition
  DQNB_TRY(dqnb_learner_write_dev(h->sub, DQNB_ONLINE, h->theta_exp, P, s));
  DQNB_TRY(dqnb_learner_write_dev(h->sub, DQNB_TARGET, static_cast<const float*>(theta), P, s));
  DQNB_TRY(dqnb_learner_update(h->sub, h->mb_s0 + M * h->img, h->mb_a + M, h->mb_r + M, h->mb_d + M,
                               h->mb_s1 + M * h->img, s));
  if (h->cfg.second_order) {
    meta_second_kernel<<<kLossBlocks, 256, 0, st>>>(static_cast<const float*>(sub_grad), h->G, h->mu1, h->nu1, h->wdir,
                                                    h->extra, h->loss_part, P, h->lcfg.learning_rate, h->lcfg.decay,
                                                    h->lcfg.eps);
    DQNB_CUDA(cudaGetLastError());
    DQNB_TRY(dqnb_learner_hvp(h->sub, h->wdir, h->hv, s));     // at theta_exp, on the online transition
    meta_second_finish_kernel<<<pb, 256, 0, st>>>(h->hv, h->G, h->mu1, h->nu1, h->extra, h->v, P,
                                                  h->lcfg.learning_rate, h->lcfg.decay, h->lcfg.eps);
  } else {
    meta_v_kernel<<<kLossBlocks, 256, 0, st>>>(static_cast<const float*>(sub_grad), h->G, h->mu1, h->nu1, h->v,
                                               h->loss_part, P, h->lcfg.learning_rate, h->lcfg.decay, h->lcfg.eps);
  }
  DQNB_TRY(dqnb_learner_adjoint_dots(h->bat, h->v, h->dot_part, s));
  logits_adam_kernel<<<1, 256, 0, st>>>(h->logits, h->p, h->dot_part, h->loss_part, h->dldp, h->adam_m, h->adam_v,
                                        h->out, h->out + M, M, h->count + 1, h->cfg.learning_rate, h->cfg.b1,
                                        h->cfg.b2, h->cfg.eps);
  DQNB_CUDA(cudaGetLastError());
  DQNB_CUDA(cudaMemcpyAsync(h->pinned + M, h->out, sizeof(float) * (M + 1), cudaMemcpyDeviceToHost, st));
  DQNB_CUDA(cudaStreamSynchronize(st));
  h->count += 1;      // the Adam state advanced only if everything above ran
  memcpy(new_logits_host, h->pinned + M, sizeof(float) * M);
  if (meta_loss_host) *meta_loss_host = h->pinned[2 * M];
  return DQNB_OK;
}

int stage_online_transition(dqnb_mgsc* h, const uint8_t* s_tm1, int32_t a, float r, float disc, const uint8_t* s_t,
                            cudaStream_t st) {
  const int M = h->M;
  // through pinned staging (a pageable source makes each copy a blocking, double-hop transfer); the
  // buffer is free again when run_meta_update returns (it synchronises the stream)
  memcpy(h->pinned_on, s_tm1, h->img);
  memcpy(h->pinned_on + h->img, s_t, h->img);
  DQNB_CUDA(cudaMemcpyAsync(h->mb_s0 + M * h->img, h->pinned_on, h->img, cudaMemcpyHostToDevice, st));
  DQNB_CUDA(cudaMemcpyAsync(h->mb_s1 + M * h->img, h->pinned_on + h->img, h->img, cudaMemcpyHostToDevice, st));
  // scalars travel by value through the pinned block (tail), then one small copy each
  int32_t* pa = reinterpret_cast<int32_t*>(h->pinned + 2 * M + 4);
  pa[0] = a;
  h->pinned[2 * M + 5] = r;
  h->pinned[2 * M + 6] = disc;
  DQNB_CUDA(cudaMemcpyAsync(h->mb_a + M, pa, 4, cudaMemcpyHostToDevice, st));
  DQNB_CUDA(cudaMemcpyAsync(h->mb_r + M, h->pinned + 2 * M + 5, 4, cudaMemcpyHostToDevice, st));
  DQNB_CUDA(cudaMemcpyAsync(h->mb_d + M, h->pinned + 2 * M + 6, 4, cudaMemcpyHostToDevice, st));
  return DQNB_OK;
}
}  // namespace

extern "C" int dqnb_mgsc_create(dqnb_learner* learner, const dqnb_mgsc_config* cfg, dqnb_mgsc** out) {
  DQNB_REQUIRE(learner && cfg && out, "null argument");
  DQNB_REQUIRE(cfg->meta_batch_size >= 1 && cfg->meta_batch_size <= kMaxMeta, "meta_batch_size %d not in 1..%d",
               cfg->meta_batch_size, kMaxMeta);
  dqnb_mgsc* h = new dqnb_mgsc();
  // everything below may fail half way: the handle is destroyed (destroy tolerates partial initialisation)
  const int rc_create = [&]() -> int {
  h->main = learner;
  h->cfg = *cfg;
  h->M = cfg->meta_batch_size;
  int rc = dqnb_learner_get_config(learner, &h->lcfg, &h->device);
  if (rc == DQNB_OK && h->lcfg.double_q) {
    set_error("MGSC uses the q_learning loss (dqn_mgsc_batched/agent.py:104-127): create the learner with double_q=0");
    rc = DQNB_EINVAL;
  }
  if (rc != DQNB_OK) return rc;
  dqnb_learner_config sc = h->lcfg;
  sc.batch_size = 1; sc.loss_mean_over = 1; sc.double_q = 0; sc.defer_apply = 1;
  rc = dqnb_learner_create(&sc, h->device, &h->sub);
  if (rc != DQNB_OK) return rc;
  sc.batch_size = h->M; sc.row_weighted = 1;
  rc = dqnb_learner_create(&sc, h->device, &h->bat);
  if (rc != DQNB_OK) return rc;
  DeviceGuard guard(h->device);
  DQNB_TRY(dqnb_learner_num_params(learner, &h->P));
  h->img = 84 * 84 * 4;
  const size_t M1 = (size_t)h->M + 1, P = (size_t)h->P;
  DQNB_TRY(dalloc(&h->mb_s0, M1 * h->img)); DQNB_TRY(dalloc(&h->mb_s1, M1 * h->img));
  DQNB_TRY(dalloc(&h->mb_a, M1)); DQNB_TRY(dalloc(&h->mb_r, M1)); DQNB_TRY(dalloc(&h->mb_d, M1));
  DQNB_TRY(dalloc(&h->G, P)); DQNB_TRY(dalloc(&h->theta_exp, P)); DQNB_TRY(dalloc(&h->mu1, P));
  DQNB_TRY(dalloc(&h->nu1, P)); DQNB_TRY(dalloc(&h->v, P));
  if (h->cfg.second_order) { DQNB_TRY(dalloc(&h->wdir, P)); DQNB_TRY(dalloc(&h->extra, P)); DQNB_TRY(dalloc(&h->hv, P)); }
  DQNB_TRY(dalloc(&h->logits, M1)); DQNB_TRY(dalloc(&h->p, M1)); DQNB_TRY(dalloc(&h->dldp, M1));
  DQNB_TRY(dalloc(&h->dot_part, M1)); DQNB_TRY(dalloc(&h->loss_part, kLossBlocks));
  DQNB_TRY(dalloc(&h->adam_m, M1)); DQNB_TRY(dalloc(&h->adam_v, M1)); DQNB_TRY(dalloc(&h->out, M1 + 1));
  DQNB_CUDA(cudaMallocHost(&h->pinned, sizeof(float) * (2 * M1 + 8)));
  DQNB_CUDA(cudaMallocHost(&h->pinned_on, 2 * h->img));
  return DQNB_OK;
  }();
  if (rc_create != DQNB_OK) { dqnb_mgsc_destroy(h); return rc_create; }
  *out = h;
  return DQNB_OK;
}

extern "C" int dqnb_mgsc_destroy(dqnb_mgsc* h) {
  if (!h) return DQNB_OK;
  DeviceGuard guard(h->device);
  cudaDeviceSynchronize();
  if (h->sub) dqnb_learner_destroy(h->sub);
  if (h->bat) dqnb_learner_destroy(h->bat);
  void* bufs[] = {h->mb_s0, h->mb_s1, h->mb_a, h->mb_r, h->mb_d, h->G, h->theta_exp, h->mu1, h->nu1, h->v,
                  h->logits, h->p, h->dldp, h->dot_part, h->loss_part, h->adam_m, h->adam_v, h->out,
                  h->wdir, h->extra, h->hv};
  for (void* b : bufs) cudaFree(b);
  if (h->pinned) cudaFreeHost(h->pinned);
  if (h->pinned_on) cudaFreeHost(h->pinned_on);
  delete h;
  return DQNB_OK;
}

extern "C" int dqnb_mgsc_meta_update_host(dqnb_mgsc* h, const uint8_t* s_tm1, const int32_t* a, const float* r,
                                          const float* disc, const uint8_t* s_t, const float* logits,
                                          const uint8_t* on_s_tm1, int32_t on_a, float on_r, float on_disc,
                                          const uint8_t* on_s_t, float* new_logits, float* meta_loss,
                                          uintptr_t stream) {
  DQNB_REQUIRE(h && s_tm1 && a && r && disc && s_t && logits && on_s_tm1 && on_s_t && new_logits, "null argument");
  DeviceGuard guard(h->device);
  cudaStream_t st = as_stream(stream);
  const size_t M = h->M;
  DQNB_CUDA(cudaMemcpyAsync(h->mb_s0, s_tm1, M * h->img, cudaMemcpyHostToDevice, st));
  DQNB_CUDA(cudaMemcpyAsync(h->mb_s1, s_t, M * h->img, cudaMemcpyHostToDevice, st));
  DQNB_CUDA(cudaMemcpyAsync(h->mb_a, a, 4 * M, cudaMemcpyHostToDevice, st));
  DQNB_CUDA(cudaMemcpyAsync(h->mb_r, r, 4 * M, cudaMemcpyHostToDevice, st));
  DQNB_CUDA(cudaMemcpyAsync(h->mb_d, disc, 4 * M, cudaMemcpyHostToDevice, st));
  DQNB_CUDA(cudaStreamSynchronize(st));   // caller buffers are pageable: do not hold on to them
  memcpy(h->pinned, logits, sizeof(float) * M);
  DQNB_TRY(stage_online_transition(h, on_s_tm1, on_a, on_r, on_disc, on_s_t, st));
  return run_meta_update(h, new_logits, meta_loss, st);
}

extern "C" int dqnb_mgsc_meta_update(dqnb_mgsc* h, dqnb_replay* replay, const int64_t* slots_host,
                                     const float* logits, const uint8_t* on_s_tm1, int32_t on_a, float on_r,
                                     float on_disc, const uint8_t* on_s_t, float* new_logits, float* meta_loss,
                                     uintptr_t stream) {
  DQNB_REQUIRE(h && replay && slots_host && logits && on_s_tm1 && on_s_t && new_logits, "null argument");
  int H, W, S;
  replay_frame_shape(replay, &H, &W, &S);
  DQNB_REQUIRE(H == 84 && W == 84 && S == 4, "replay frames %dx%dx%d are not 84x84x4", H, W, S);
  DeviceGuard guard(h->device);
  cudaStream_t st = as_stream(stream);
  DQNB_TRY(replay_gather_into(replay, slots_host, h->M, h->mb_s0, h->mb_s1, h->mb_a, h->mb_r, h->mb_d, st));
  memcpy(h->pinned, logits, sizeof(float) * h->M);
  DQNB_TRY(stage_online_transition(h, on_s_tm1, on_a, on_r, on_disc, on_s_t, st));
  return run_meta_update(h, new_logits, meta_loss, st);
}

namespace {
int mgsc_locate(dqnb_mgsc* h, int which, float** p, int64_t* n) {
  switch (which) {
    case DQNB_MGSC_ADAM_M: *p = h->adam_m; *n = h->M; break;
    case DQNB_MGSC_ADAM_V: *p = h->adam_v; *n = h->M; break;
    case DQNB_MGSC_P: *p = h->p; *n = h->M; break;
    case DQNB_MGSC_DL_DP: *p = h->dldp; *n = h->M; break;
    case DQNB_MGSC_G: *p = h->G; *n = h->P; break;
    case DQNB_MGSC_V: *p = h->v; *n = h->P; break;
    case DQNB_MGSC_THETA_EXP: *p = h->theta_exp; *n = h->P; break;
    default: set_error("unknown MGSC array %d", which); return DQNB_EINVAL;
  }
  return DQNB_OK;
}
}  // namespace

extern "C" int dqnb_mgsc_read(dqnb_mgsc* h, int which, float* host, int64_t n) {
  DQNB_REQUIRE(h && host, "null argument");
  DeviceGuard guard(h->device);
  float* p; int64_t len;
  DQNB_TRY(mgsc_locate(h, which, &p, &len));
  DQNB_REQUIRE(n == len, "length %lld != %lld", (long long)n, (long long)len);
  DQNB_CUDA(cudaMemcpy(host, p, sizeof(float) * n, cudaMemcpyDeviceToHost));
  return DQNB_OK;
}

// Test hook: a state array (or debug tap) of the internal batch-of-M (0) / batch-of-1 (1) learner as the
// last meta_update left it - the parity tests take the device's own ReLU masks and td errors from here.
extern "C" int dqnb_mgsc_read_learner(dqnb_mgsc* h, int which_learner, int which, float* host, int64_t n) {
  DQNB_REQUIRE(h && host, "null argument");
  DQNB_REQUIRE(which_learner == 0 || which_learner == 1, "which_learner must be 0 (meta batch) or 1 (online transition)");
  return dqnb_learner_read(which_learner == 0 ? h->bat : h->sub, which, host, n, 0);
}

extern "C" int dqnb_mgsc_write(dqnb_mgsc* h, int which, const float* host, int64_t n) {
  DQNB_REQUIRE(h && host, "null argument");
  DQNB_REQUIRE(which == DQNB_MGSC_ADAM_M || which == DQNB_MGSC_ADAM_V, "MGSC array %d is read-only", which);
  DeviceGuard guard(h->device);
  float* p; int64_t len;
  DQNB_TRY(mgsc_locate(h, which, &p, &len));
  DQNB_REQUIRE(n == len, "length %lld != %lld", (long long)n, (long long)len);
  DQNB_CUDA(cudaMemcpy(p, host, sizeof(float) * n, cudaMemcpyHostToDevice));
  return DQNB_OK;
}

extern "C" int dqnb_mgsc_adam_count(dqnb_mgsc* h, int32_t* get, const int32_t* set) {
  DQNB_REQUIRE(h, "null argument");
  if (set) h->count = *set;
  if (get) *get = h->count;
  return DQNB_OK;
}
