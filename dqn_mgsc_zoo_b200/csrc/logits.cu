// K8a: the MGSC logit ring in HBM (replay_circular.py:148-248 CircularLogitBuffer and the logits
// half of MGSCReservoirDistribution :500-560).  The reference keeps a float32 NumPy array of
// `capacity` logits (-inf = empty slot) on the host and walks all of it on every add
// (logsumexp, :166-179) and every sample (softmax + float64 cumsum + searchsorted inside
// Generator.choice, :205-217): 1.94 ms per add and 57 ms per sample at capacity 1e6.
// Here the array lives on the device (4 MB at 1e6: L2-resident), head/tail/size bookkeeping and
// the random draws stay on the host (same Generator stream as the reference):
//   fill   : logits[i] = logsumexp(all) - log(size)          two small kernels, no host round trip
//   sample : p = exp(l - lse) in float32, CDF in float64 (blocked prefix sums + in-order scan inside
//            the hit block), cdf/cdf[-1] <= u  counted like searchsorted(side='right')
// HBM/L2-bound streaming reductions; no tensor cores.
// Running logsumexp: walking the whole ring for every add is the reference's 1.94 ms hotspot and was still three
// passes over 4 MB here.  The device keeps (c, S) with S = sum_i exp(l_i - c) in float64, c >= every logit written
// since the last exact pass; fill / set / popleft adjust S by the terms they add and remove (single-CTA kernels,
// fixed summation order), so the new logit lse - log(size) costs one tiny launch.  An exact pass (the kernels
// below) re-anchors (c, S) every kExactEvery mutations and after any whole-array write, so float64 drift stays
// far below the float32 rounding of the logits themselves (the tests hold them to 2e-6 of the reference).
#include <math.h>
#include <stdlib.h>

#include <algorithm>
#include <vector>

#include "common.cuh"

using namespace dqnb;

namespace {
constexpr int kBlk = 1024;        // logits per CDF block
constexpr int kThreads = 256;

__device__ __forceinline__ float block_max(float v, float* red) {
  const int t = threadIdx.x;
  red[t] = v;
  __syncthreads();
  for (int s = kThreads / 2; s > 0; s >>= 1) { if (t < s) red[t] = fmaxf(red[t], red[t + s]); __syncthreads(); }
  const float r = red[0];
  __syncthreads();
  return r;
}
template <class T>
__device__ __forceinline__ T block_sum(T v, T* red) {
  const int t = threadIdx.x;
  red[t] = v;
  __syncthreads();
  for (int s = kThreads / 2; s > 0; s >>= 1) { if (t < s) red[t] += red[t + s]; __syncthreads(); }
  const T r = red[0];
  __syncthreads();
  return r;
}

// pass 1: per-block max
__global__ void max_kernel(const float* __restrict__ l, int64_t n, float* __restrict__ part) {
  __shared__ float red[kThreads];
  float m = -INFINITY;
  for (int64_t i = (int64_t)blockIdx.x * kBlk + threadIdx.x; i < min((int64_t)(blockIdx.x + 1) * kBlk, n); i += kThreads)
    m = fmaxf(m, l[i]);
  m = block_max(m, red);
  if (threadIdx.x == 0) part[blockIdx.x] = m;
}
// pass 2 (one block): global max c -> stat[0]
__global__ void max_final_kernel(const float* __restrict__ part, int nb, float* __restrict__ stat) {
  __shared__ float red[kThreads];
  float m = -INFINITY;
  for (int i = threadIdx.x; i < nb; i += kThreads) m = fmaxf(m, part[i]);
  m = block_max(m, red);
  if (threadIdx.x == 0) stat[0] = m;
}
// pass 3: per-block sum of exp(l - c) in float32 (the reference sums float32, np.sum pairwise)
__global__ void sumexp_kernel(const float* __restrict__ l, int64_t n, const float* __restrict__ stat,
                              float* __restrict__ part) {
  __shared__ float red[kThreads];
  const float c = stat[0];
  float s = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * kBlk + threadIdx.x; i < min((int64_t)(blockIdx.x + 1) * kBlk, n); i += kThreads)
    s += expf(l[i] - c);
  s = block_sum(s, red);
  if (threadIdx.x == 0) part[blockIdx.x] = s;
}
// pass 4 (one block): lse = c + log(sum) -> stat[1]; optionally logits[fill_at] = lse - log(size)
__global__ void lse_final_kernel(const float* __restrict__ part, int nb, float* __restrict__ stat,
                                 float* __restrict__ l, int64_t fill_at, double log_size, double* __restrict__ run) {
  __shared__ float red[kThreads];
  float s = 0.f;
  for (int i = threadIdx.x; i < nb; i += kThreads) s += part[i];
  s = block_sum(s, red);
  if (threadIdx.x == 0) {
    const float lse = stat[0] + logf(s);
    stat[1] = lse;
    stat[2] = (s != s) ? 1.f : 0.f;   // a NaN logit (fmaxf drops it from the max, expf carries it into the sum)
    // np.float32 lse minus np.log(int) (float64), rounded when stored into the float32 array
    double c = (double)stat[0], S = (double)s;
    if (fill_at >= 0) {
      const float old = l[fill_at], nv = (float)((double)lse - log_size);
      l[fill_at] = nv;
      if (nv > stat[0]) { S *= exp(c - (double)nv); c = (double)nv; }
      S += (double)expf(nv - (float)c) - (double)expf(old - (float)c);
    }
    run[0] = c; run[1] = S;      // anchor of the running logsumexp
  }
}
// Running-mode twins (one CTA).  run = {c, S}.
// fill: l[at] = (float)(c + log S) - log(size), S adjusted for the slot's old and new value.
__global__ void run_fill_kernel(float* __restrict__ l, int64_t at, double log_size, double* __restrict__ run,
                                float* __restrict__ stat) {
  if (threadIdx.x != 0) return;
  double c = run[0], S = run[1];
  const float lse = (float)(c + log(S));
  const float old = l[at], nv = (float)((double)lse - log_size);
  l[at] = nv;
  if ((double)nv > c) { S *= exp(c - (double)nv); c = (double)nv; }
  S += (double)expf(nv - (float)c) - (double)expf(old - (float)c);
  run[0] = c; run[1] = S;
  stat[1] = lse;
}
// set: l[idx[k]] = v[k] for k < n (indices distinct), S adjusted; idx == nullptr: the by-value case n == 1.
__global__ void __launch_bounds__(kThreads) run_set_kernel(float* __restrict__ l, const int64_t* __restrict__ idx,
                                                           const float* __restrict__ v, int n, int64_t idx0, float v0,
                                                           double* __restrict__ run) {
  __shared__ float redf[kThreads];
  __shared__ double redd[kThreads];
  __shared__ double c_s;
  const int t = threadIdx.x;
  float vmax = -INFINITY;
  for (int k = t; k < n; k += kThreads) vmax = fmaxf(vmax, idx ? v[k] : v0);
  vmax = block_max(vmax, redf);
  if (t == 0) {
    double c = run[0];
    if ((double)vmax > c) { run[1] *= exp(c - (double)vmax); c = (double)vmax; run[0] = c; }
    c_s = c;
  }
  __syncthreads();
  const float c = (float)c_s;
  double delta = 0.0;
  for (int k = t; k < n; k += kThreads) {
    const int64_t i = idx ? idx[k] : idx0;
    const float nv = idx ? v[k] : v0, old = l[i];
    l[i] = nv;
    delta += (double)expf(nv - c) - (double)expf(old - c);
  }
  delta = block_sum(delta, redd);
  if (t == 0) run[1] += delta;
}
// stat[1] = float32 lse from the running state (sample's softmax reference point)
__global__ void run_lse_kernel(const double* __restrict__ run, float* __restrict__ stat) {
  if (threadIdx.x == 0) { stat[1] = (float)(run[0] + log(run[1])); stat[2] = 0.f; }
}
// CDF pass: block sums of p = exp(l - lse) accumulated in float64, in index order inside a thread's
// contiguous run so the blocked sum stays close to the sequential cumsum of the reference.
__global__ void psum_kernel(const float* __restrict__ l, int64_t n, const float* __restrict__ stat,
                            double* __restrict__ part) {
  __shared__ double red[kThreads];
  const float lse = stat[1];
  constexpr int RUN = kBlk / kThreads;
  const int64_t beg = (int64_t)blockIdx.x * kBlk + threadIdx.x * RUN;
  double s = 0.0;
#pragma unroll
  for (int k = 0; k < RUN; ++k)
    if (beg + k < n) s += (double)expf(l[beg + k] - lse);
  s = block_sum(s, red);
  if (threadIdx.x == 0) part[blockIdx.x] = s;
}
// (one warp) inclusive prefix of the block sums: lane-contiguous chunks summed in order, a warp
// scan of the chunk totals, then the chunk re-walked with its offset.
__global__ void pscan_kernel(double* __restrict__ part, int nb, double* __restrict__ total) {
  const int lane = threadIdx.x;
  const int chunk = (nb + 31) / 32, beg = lane * chunk, end = min(beg + chunk, nb);
  double s = 0.0;
  for (int i = beg; i < end; ++i) s += part[i];
  double incl = s;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const double t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  double acc = incl - s;
  for (int i = beg; i < end; ++i) { acc += part[i]; part[i] = acc; }
  if (lane == 31) *total = incl;
}
// One warp per uniform: binary search over the block prefixes, then the hit block (kBlk logits) is
// scanned by the warp - lane j sums its 32 consecutive entries in index order, a warp scan of the lane
// totals gives every lane its offset, and each lane looks for the first entry whose cdf / total > x
// (searchsorted(side='right') counts the entries <= x).  The exact float64 division is only evaluated
// when the product test is within rounding of the boundary.
__device__ __forceinline__ bool cdf_above(double acc, double tot, double x) {
  const double t = x * tot;
  if (acc > t * (1.0 + 1e-12)) return true;
  if (acc < t * (1.0 - 1e-12)) return false;
  return acc / tot > x;
}
__global__ void search_kernel(const float* __restrict__ l, int64_t n, const float* __restrict__ stat,
                              const double* __restrict__ prefix, int nb, const double* __restrict__ total,
                              const double* __restrict__ u, int count, int64_t* __restrict__ out) {
  const int k = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (k >= count) return;
  const double tot = *total, x = u[k];
  const float lse = stat[1];
  // first block whose inclusive prefix / total > x
  int lo = 0, hi = nb - 1;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (prefix[mid] / tot > x) hi = mid; else lo = mid + 1;
  }
  const double base = lo > 0 ? prefix[lo - 1] : 0.0;
  constexpr int RUN = kBlk / 32;
  const int64_t beg = (int64_t)lo * kBlk + lane * RUN, end = min((int64_t)(lo + 1) * kBlk, n);
  double p[RUN], run = 0.0;
#pragma unroll
  for (int i = 0; i < RUN; ++i) {
    p[i] = beg + i < end ? (double)expf(l[beg + i] - lse) : 0.0;
    run += p[i];
  }
  double incl = run;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const double t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  double acc = base + (incl - run);
  int64_t idx = -1;
#pragma unroll
  for (int i = 0; i < RUN; ++i) {
    acc += p[i];
    if (idx < 0 && beg + i < end && cdf_above(acc, tot, x)) idx = beg + i;
  }
  const unsigned hit = __ballot_sync(0xffffffffu, idx >= 0);
  if (hit != 0u) {
    idx = __shfl_sync(0xffffffffu, idx, __ffs(hit) - 1);
  } else if (lane == 0) {
    // The block's own re-summation came out a last-ulp short of the prefix that selected it.  searchsorted on
    // the reference's cdf then lands on the next entry that carries probability (the cdf is flat over empty
    // -inf slots, which it can never return); past the end of the array, on the last such entry.
    idx = end;
    while (idx < n && !(expf(l[idx] - lse) > 0.f)) ++idx;
    if (idx >= n) { idx = n - 1; while (idx > 0 && !(expf(l[idx] - lse) > 0.f)) --idx; }
  }
  if (lane == 0) out[k] = idx;
}
__global__ void set_kernel(float* __restrict__ l, const int64_t* __restrict__ idx, const float* __restrict__ v, int n) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) l[idx[k]] = v[k];
}
__global__ void get_kernel(const float* __restrict__ l, const int64_t* __restrict__ idx, float* __restrict__ v, int n) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) v[k] = l[idx[k]];
}
__global__ void fill_value_kernel(float* __restrict__ l, int64_t n, float v) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) l[i] = v;
}
}  // namespace

struct dqnb_logits {
  int device = 0;
  int64_t capacity = 0;
  int nb = 0;
  float* l = nullptr;
  float* part = nullptr;      // [nb]
  float* stat = nullptr;      // [0] max, [1] lse
  double* dpart = nullptr;    // [nb] + [1] total
  double* run = nullptr;      // {c, S}: running logsumexp anchor (device)
  bool run_valid = false;     // (c, S) describe the current array
  int run_ops = 0;            // mutations applied to (c, S) since the last exact pass
  // small pinned/device scratch for index lists
  int64_t io_cap = 0;
  int64_t* d_idx = nullptr;
  float* d_val = nullptr;
  double* d_u = nullptr;
  uint8_t* pinned = nullptr;  // io_cap * 16 bytes
};

namespace {
constexpr int kExactEvery = 64;     // running-mode mutations between two exact passes
const bool kRunningLse = getenv("DQNB_NO_RUNNING_LSE") == nullptr;
int lse_passes(dqnb_logits* h, int64_t fill_at, double log_size, cudaStream_t st) {
  max_kernel<<<h->nb, kThreads, 0, st>>>(h->l, h->capacity, h->part);
  max_final_kernel<<<1, kThreads, 0, st>>>(h->part, h->nb, h->stat);
  sumexp_kernel<<<h->nb, kThreads, 0, st>>>(h->l, h->capacity, h->stat, h->part);
  lse_final_kernel<<<1, kThreads, 0, st>>>(h->part, h->nb, h->stat, h->l, fill_at, log_size, h->run);
  DQNB_CUDA(cudaGetLastError());
  h->run_valid = kRunningLse;
  h->run_ops = 0;
  return DQNB_OK;
}
// true when the next operation may use (c, S) instead of walking the array
bool running(dqnb_logits* h) {
  if (h->run_valid && h->run_ops >= kExactEvery) h->run_valid = false;
  return h->run_valid;
}
int ensure_io(dqnb_logits* h, int64_t n) {
  if (n <= h->io_cap) return DQNB_OK;
  DQNB_CUDA(cudaDeviceSynchronize());
  cudaFree(h->d_idx); cudaFree(h->d_val); cudaFree(h->d_u);
  if (h->pinned) cudaFreeHost(h->pinned);
  h->io_cap = std::max<int64_t>(n, 1024);
  DQNB_CUDA(cudaMalloc(&h->d_idx, sizeof(int64_t) * h->io_cap));
  DQNB_CUDA(cudaMalloc(&h->d_val, sizeof(float) * h->io_cap));
  DQNB_CUDA(cudaMalloc(&h->d_u, sizeof(double) * h->io_cap));
  DQNB_CUDA(cudaMallocHost(&h->pinned, 16 * (size_t)h->io_cap));
  return DQNB_OK;
}
int check_indices(dqnb_logits* h, const int64_t* idx, int64_t n) {
  for (int64_t i = 0; i < n; ++i)
    DQNB_REQUIRE(idx[i] >= 0 && idx[i] < h->capacity, "logit index %lld outside [0,%lld)", (long long)idx[i],
                 (long long)h->capacity);
  return DQNB_OK;
}
}  // namespace

extern "C" int dqnb_logits_create(int64_t capacity, int device, dqnb_logits** out) {
  DQNB_REQUIRE(out && capacity >= 1 && capacity <= (int64_t)1 << 31, "bad capacity %lld", (long long)capacity);
  DeviceGuard guard(device);
  dqnb_logits* h = new dqnb_logits();
  // everything below may fail half way: the handle is destroyed (destroy tolerates partial initialisation)
  const int rc_create = [&]() -> int {
  h->device = device; h->capacity = capacity;
  h->nb = (int)((capacity + kBlk - 1) / kBlk);
  DQNB_CUDA(cudaMalloc(&h->l, sizeof(float) * capacity));
  DQNB_CUDA(cudaMalloc(&h->part, sizeof(float) * h->nb));
  DQNB_CUDA(cudaMalloc(&h->stat, sizeof(float) * 4));
  DQNB_CUDA(cudaMalloc(&h->dpart, sizeof(double) * (h->nb + 1)));
  DQNB_CUDA(cudaMalloc(&h->run, sizeof(double) * 2));
  fill_value_kernel<<<(unsigned)((capacity + 255) / 256), 256>>>(h->l, capacity, -INFINITY);
  DQNB_CUDA(cudaGetLastError());
  DQNB_CUDA(cudaDeviceSynchronize());
  return DQNB_OK;
  }();
  if (rc_create != DQNB_OK) { dqnb_logits_destroy(h); return rc_create; }
  *out = h;
  return DQNB_OK;
}

extern "C" int dqnb_logits_destroy(dqnb_logits* h) {
  if (!h) return DQNB_OK;
  DeviceGuard guard(h->device);
  cudaDeviceSynchronize();
  cudaFree(h->l); cudaFree(h->part); cudaFree(h->stat); cudaFree(h->dpart); cudaFree(h->run);
  cudaFree(h->d_idx); cudaFree(h->d_val); cudaFree(h->d_u);
  if (h->pinned) cudaFreeHost(h->pinned);
  delete h;
  return DQNB_OK;
}

// logits[index] = size == 0 ? 0 : logsumexp(all) - log(size)   (replay_circular.py:166-179, :526-533)
extern "C" int dqnb_logits_fill(dqnb_logits* h, int64_t index, int64_t size, uintptr_t stream) {
  DQNB_REQUIRE(h && index >= 0 && index < h->capacity && size >= 0, "bad argument");
  DeviceGuard guard(h->device);
  cudaStream_t st = as_stream(stream);
  if (size == 0) {
    fill_value_kernel<<<1, 32, 0, st>>>(h->l + index, 1, 0.0f);
    DQNB_CUDA(cudaGetLastError());
    h->run_valid = false;
    return DQNB_OK;
  }
  if (running(h)) {
    run_fill_kernel<<<1, 32, 0, st>>>(h->l, index, log((double)size), h->run, h->stat);
    DQNB_CUDA(cudaGetLastError());
    h->run_ops++;
    return DQNB_OK;
  }
  return lse_passes(h, index, log((double)size), st);
}

extern "C" int dqnb_logits_set(dqnb_logits* h, const int64_t* indices, const float* values, int64_t n,
                               uintptr_t stream) {
  DQNB_REQUIRE(h && (n == 0 || (indices && values)), "null argument");
  if (n == 0) return DQNB_OK;
  DQNB_TRY(check_indices(h, indices, n));
  DeviceGuard guard(h->device);
  cudaStream_t st = as_stream(stream);
  for (int64_t i = 0; i < n; ++i)
    if (values[i] != values[i] || values[i] == INFINITY) h->run_valid = false;   // NaN / +inf: only the exact pass handles them
  if (n == 1) {   // the per-add case (popleft / explicit priority): by value, nothing to wait for
    if (running(h)) {
      run_set_kernel<<<1, kThreads, 0, st>>>(h->l, nullptr, nullptr, 1, indices[0], values[0], h->run);
      h->run_ops++;
    } else {
      fill_value_kernel<<<1, 32, 0, st>>>(h->l + indices[0], 1, values[0]);
    }
    DQNB_CUDA(cudaGetLastError());
    return DQNB_OK;
  }
  DQNB_TRY(ensure_io(h, n));
  DQNB_CUDA(cudaStreamSynchronize(st));   // pinned scratch is reused call to call
  int64_t* pi = reinterpret_cast<int64_t*>(h->pinned);
  float* pv = reinterpret_cast<float*>(h->pinned + 8 * (size_t)h->io_cap);
  // duplicate indices: NumPy's `l[idx] = v` keeps the last one; the running kernel needs them distinct
  int64_t m = 0;
  if (n <= 4096) {
    for (int64_t i = 0; i < n; ++i) {
      bool later = false;
      for (int64_t j = i + 1; j < n && !later; ++j) later = indices[j] == indices[i];
      if (!later) { pi[m] = indices[i]; pv[m] = values[i]; ++m; }
    }
  } else {      // big scatter (whole-array style use): plain kernel, exact pass next time
    memcpy(pi, indices, sizeof(int64_t) * n);
    memcpy(pv, values, sizeof(float) * n);
    m = n;
    h->run_valid = false;
  }
  DQNB_CUDA(cudaMemcpyAsync(h->d_idx, pi, sizeof(int64_t) * m, cudaMemcpyHostToDevice, st));
  DQNB_CUDA(cudaMemcpyAsync(h->d_val, pv, sizeof(float) * m, cudaMemcpyHostToDevice, st));
  if (running(h)) {
    run_set_kernel<<<1, kThreads, 0, st>>>(h->l, h->d_idx, h->d_val, (int)m, 0, 0.f, h->run);
    h->run_ops++;
  } else {
    set_kernel<<<(unsigned)((m + 255) / 256), 256, 0, st>>>(h->l, h->d_idx, h->d_val, (int)m);
  }
  DQNB_CUDA(cudaGetLastError());
  return DQNB_OK;
}

extern "C" int dqnb_logits_get(dqnb_logits* h, const int64_t* indices, float* values, int64_t n, uintptr_t stream) {
  DQNB_REQUIRE(h && (n == 0 || (indices && values)), "null argument");
  if (n == 0) return DQNB_OK;
  DQNB_TRY(check_indices(h, indices, n));
  DeviceGuard guard(h->device);
  cudaStream_t st = as_stream(stream);
  DQNB_TRY(ensure_io(h, n));
  DQNB_CUDA(cudaStreamSynchronize(st));
  int64_t* pi = reinterpret_cast<int64_t*>(h->pinned);
  float* pv = reinterpret_cast<float*>(h->pinned + 8 * (size_t)h->io_cap);
  memcpy(pi, indices, sizeof(int64_t) * n);
  DQNB_CUDA(cudaMemcpyAsync(h->d_idx, pi, sizeof(int64_t) * n, cudaMemcpyHostToDevice, st));
  get_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(h->l, h->d_idx, h->d_val, (int)n);
  DQNB_CUDA(cudaGetLastError());
  DQNB_CUDA(cudaMemcpyAsync(pv, h->d_val, sizeof(float) * n, cudaMemcpyDeviceToHost, st));
  DQNB_CUDA(cudaStreamSynchronize(st));
  memcpy(values, pv, sizeof(float) * n);
  return DQNB_OK;
}

// Generator.choice(capacity, n, p=softmax(logits)) given the n doubles Generator.random(n) drew
// (replay_circular.py:205-217); returns absolute slot indices.
extern "C" int dqnb_logits_sample(dqnb_logits* h, const double* uniforms, int64_t n, int64_t* indices,
                                  uintptr_t stream) {
  DQNB_REQUIRE(h && (n == 0 || (uniforms && indices)), "null argument");
  if (n == 0) return DQNB_OK;
  DeviceGuard guard(h->device);
  cudaStream_t st = as_stream(stream);
  DQNB_TRY(ensure_io(h, n));
  DQNB_CUDA(cudaStreamSynchronize(st));
  double* pu = reinterpret_cast<double*>(h->pinned);
  memcpy(pu, uniforms, sizeof(double) * n);
  DQNB_CUDA(cudaMemcpyAsync(h->d_u, pu, sizeof(double) * n, cudaMemcpyHostToDevice, st));
  if (running(h)) run_lse_kernel<<<1, 32, 0, st>>>(h->run, h->stat);
  else DQNB_TRY(lse_passes(h, -1, 0.0, st));
  psum_kernel<<<h->nb, kThreads, 0, st>>>(h->l, h->capacity, h->stat, h->dpart);
  pscan_kernel<<<1, 32, 0, st>>>(h->dpart, h->nb, h->dpart + h->nb);
  search_kernel<<<(unsigned)((n + 1) / 2), 64, 0, st>>>(h->l, h->capacity, h->stat, h->dpart, h->nb,
                                                         h->dpart + h->nb, h->d_u, (int)n, h->d_idx);
  DQNB_CUDA(cudaGetLastError());
  int64_t* pi = reinterpret_cast<int64_t*>(h->pinned + 8 * (size_t)h->io_cap);
  DQNB_CUDA(cudaMemcpyAsync(pi, h->d_idx, sizeof(int64_t) * n, cudaMemcpyDeviceToHost, st));
  float* pflag = reinterpret_cast<float*>(h->pinned);   // the uniforms were consumed by the copy above
  DQNB_CUDA(cudaMemcpyAsync(pflag, h->stat + 2, sizeof(float), cudaMemcpyDeviceToHost, st));
  DQNB_CUDA(cudaStreamSynchronize(st));
  // Generator.choice raises ValueError('probabilities contain NaN') in the reference (replay_circular.py:205-217)
  DQNB_REQUIRE(*pflag == 0.f, "probabilities contain NaN");
  memcpy(indices, pi, sizeof(int64_t) * n);
  return DQNB_OK;
}

// whole-array host copies (get_state / set_state / experiments poking `_logits`); synchronous.
extern "C" int dqnb_logits_read(dqnb_logits* h, float* host, int64_t n) {
  DQNB_REQUIRE(h && host && n == h->capacity, "bad argument");
  DeviceGuard guard(h->device);
  DQNB_CUDA(cudaMemcpy(host, h->l, sizeof(float) * n, cudaMemcpyDeviceToHost));
  return DQNB_OK;
}
extern "C" int dqnb_logits_write(dqnb_logits* h, const float* host, int64_t n) {
  DQNB_REQUIRE(h && host && n == h->capacity, "bad argument");
  DeviceGuard guard(h->device);
  DQNB_CUDA(cudaMemcpy(h->l, host, sizeof(float) * n, cudaMemcpyHostToDevice));
  h->run_valid = false;
  return DQNB_OK;
}
extern "C" int dqnb_logits_lse(dqnb_logits* h, float* lse, uintptr_t stream) {
  DQNB_REQUIRE(h && lse, "null argument");
  DeviceGuard guard(h->device);
  cudaStream_t st = as_stream(stream);
  DQNB_TRY(lse_passes(h, -1, 0.0, st));
  float stat[2];
  DQNB_CUDA(cudaMemcpyAsync(stat, h->stat, sizeof(stat), cudaMemcpyDeviceToHost, st));
  DQNB_CUDA(cudaStreamSynchronize(st));
  *lse = stat[1];
  return DQNB_OK;
}

// ------------------------------------------------------------------------------------------------------------
// Allocation-free logsumexp (no counterpart in the reference).
// Same kernels as dqnb_logits_lse (max -> global max -> sum of exp -> lse), but on a caller-owned device array
// and with the scratch those passes need - per-block partials [nb], the {max, lse, nan-flag} cell, the running
// anchor - in ONE file-scope device array instead of buffers a handle allocates.  For callers that cannot call
// cudaMalloc: code under CUDA graph capture, tight serving loops, harnesses that forbid it.
//   * limit: kNoAllocMaxLogits = 4 194 304 logits (4 096 blocks of 1 024) -> 16 KB of partials + a 256-byte
//     header; every GPU that loads the library holds these 16.6 KB until the process exits;
//   * the array is SHARED by every call and keeps the last call's contents.  Nothing in it has to start at zero:
//     each pass overwrites exactly the cells the next pass reads (part[0..nb), stat[0..3), run[0..2)), so stale
//     partials of a longer earlier call are never read.  Calls that overlap on the GPU (different streams, or
//     different host threads without their own ordering) race on it: serialise them;
//   * in host code `g_noalloc_scratch` is a host-side stand-in, not a device address: the address the kernels get
//     comes from cudaGetSymbolAddress, looked up once per device with that device current.
namespace {
constexpr int kNoAllocMaxBlocks = 4096;
constexpr int64_t kNoAllocMaxLogits = (int64_t)kNoAllocMaxBlocks * kBlk;
constexpr size_t kNoAllocHeader = 256;     // run[2] doubles at 0, stat[4] floats at 64
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocHeader + sizeof(float) * kNoAllocMaxBlocks];

constexpr int kNoAllocMaxDevices = 64;
unsigned char* g_noalloc_ptr[kNoAllocMaxDevices] = {};     // device address of the array, per device ordinal

__global__ void noalloc_publish_kernel(const float* __restrict__ stat, float* __restrict__ out) {
  if (threadIdx.x == 0) *out = stat[1];
}

// device address of the scratch array on `device` (which must be current)
int noalloc_scratch(int device, unsigned char** ptr) {
  DQNB_REQUIRE(device >= 0 && device < kNoAllocMaxDevices, "device ordinal %d outside [0,%d)", device, kNoAllocMaxDevices);
  if (g_noalloc_ptr[device] == nullptr) {
    unsigned char* p = nullptr;
    DQNB_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
    DQNB_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a null scratch address");
    g_noalloc_ptr[device] = p;
  }
  *ptr = g_noalloc_ptr[device];
  return DQNB_OK;
}
}  // namespace

extern "C" int dqnb_logsumexp_noalloc(int device, const float* logits_dev, int64_t n, float* lse_dev, uintptr_t stream) {
  DQNB_REQUIRE(logits_dev && lse_dev, "null argument");
  DQNB_REQUIRE(n >= 1, "n=%lld: need at least one logit", (long long)n);
  DQNB_REQUIRE(n <= kNoAllocMaxLogits,
               "n=%lld exceeds the allocation-free limit of %lld logits (fixed scratch of %d partials); use "
               "dqnb_logits_create / dqnb_logits_lse for larger arrays",
               (long long)n, (long long)kNoAllocMaxLogits, kNoAllocMaxBlocks);
  DeviceGuard guard(device);
  unsigned char* base = nullptr;
  DQNB_TRY(noalloc_scratch(device, &base));
  double* run = reinterpret_cast<double*>(base);
  float* stat = reinterpret_cast<float*>(base + 64);
  float* part = reinterpret_cast<float*>(base + kNoAllocHeader);
  cudaStream_t st = as_stream(stream);
  const int nb = (int)((n + kBlk - 1) / kBlk);
  max_kernel<<<nb, kThreads, 0, st>>>(logits_dev, n, part);
  max_final_kernel<<<1, kThreads, 0, st>>>(part, nb, stat);
  sumexp_kernel<<<nb, kThreads, 0, st>>>(logits_dev, n, stat, part);
  lse_final_kernel<<<1, kThreads, 0, st>>>(part, nb, stat, nullptr, -1, 0.0, run);
  noalloc_publish_kernel<<<1, 32, 0, st>>>(stat, lse_dev);
  DQNB_CUDA(cudaGetLastError());
  return DQNB_OK;
}
