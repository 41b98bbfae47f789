// Prioritized replay's sum tree in device memory: replaces the float64 NumPy array + Python
// loops of SumTree (replay.py:379-559; the replay_circular.py:710-890 copy is identical).  Same layout as the
// reference - storage[2 * first_leaf] doubles, node i = storage[2i] + storage[2i+1], leaves from first_leaf,
// storage[1] the root - and the same IEEE float64 additions / subtractions in the same tree order, so every node,
// every `target -= left_sum` of a query and therefore every sampled index is bit-identical to the reference's:
//   set   (replay.py:402-414): write the leaves, then level by level recompute the touched ancestors
//         (the reference walks each index to the root one after the other; a node's final value only depends on
//         its children's final values, so the level-synchronous sweep ends in the same state);
//   query (replay.py:424-440,536-559): one thread per target descends the tree; the targets are either given
//         (SumTree.query) or formed on the device as uniform * root (PrioritizedDistribution.sample, :697-699).
// A launch touches ~40 nodes per index: latency-bound, no tensor cores, nothing to stage.
#include <algorithm>
#include <vector>

#include "common.cuh"

using namespace dqnb;

namespace {
constexpr int kSetThreads = 256;

__global__ void __launch_bounds__(kSetThreads) sumtree_set_kernel(double* __restrict__ st, int64_t first_leaf, int levels,
                                                                  const int64_t* __restrict__ idx,
                                                                  const double* __restrict__ val, int n) {
  for (int k = threadIdx.x; k < n; k += kSetThreads) st[first_leaf + idx[k]] = val[k];
  __syncthreads();
  for (int lvl = 1; lvl <= levels; ++lvl) {
    for (int k = threadIdx.x; k < n; k += kSetThreads) {
      const int64_t p = (first_leaf + idx[k]) >> lvl;      // lvl = levels reaches node 1 (the root)
      // duplicates (two indices under one parent) compute and store the same bits
      st[p] = __dadd_rn(st[2 * p], st[2 * p + 1]);
    }
    __syncthreads();
  }
}

// flag[0] |= 1 when a target is outside [0, root): the reference raises ValueError there (replay.py:538-539).
__global__ void sumtree_query_kernel(const double* __restrict__ st, int64_t first_leaf, const double* __restrict__ in,
                                     int n, int scale_by_root, int64_t* __restrict__ out, double* __restrict__ root_out,
                                     int* __restrict__ flag) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  const double root = st[1];
  if (k == 0) *root_out = root;
  if (k >= n) return;
  double target = scale_by_root ? __dmul_rn(in[k], root) : in[k];
  if (!(target >= 0.0 && target < root)) { atomicOr(flag, 1); out[k] = 0; return; }
  int64_t i = 1;
  while (i < first_leaf) {
    const double left = st[2 * i];
    if (target < left) {
      i = 2 * i;
    } else {
      i = 2 * i + 1;
      target = __dsub_rn(target, left);
    }
  }
  out[k] = i - first_leaf;
}
}  // namespace

struct dqnb_sumtree {
  int device = 0;
  int64_t first_leaf = 0;
  int levels = 0;
  double* st = nullptr;
  // staging: a small ring of pinned blocks so that `set` never waits for the previous launch
  static constexpr int kRing = 4;
  int64_t io_cap = 0;
  uint8_t* pinned[kRing] = {};
  uint8_t* dev[kRing] = {};
  cudaEvent_t ev[kRing] = {};
  bool busy[kRing] = {};
  int cur = 0;
  int* d_flag = nullptr;
};

namespace {
int ensure_io(dqnb_sumtree* h, int64_t n) {
  if (n <= h->io_cap) return DQNB_OK;
  DQNB_CUDA(cudaDeviceSynchronize());
  h->io_cap = std::max<int64_t>(n, 256);
  for (int k = 0; k < dqnb_sumtree::kRing; ++k) {
    if (h->pinned[k]) cudaFreeHost(h->pinned[k]);
    cudaFree(h->dev[k]);
    DQNB_CUDA(cudaMallocHost(&h->pinned[k], 16 * (size_t)h->io_cap + 32));
    DQNB_CUDA(cudaMalloc(&h->dev[k], 16 * (size_t)h->io_cap + 32));
    h->busy[k] = false;
  }
  return DQNB_OK;
}
int take_block(dqnb_sumtree* h, int64_t n, int* out) {
  DQNB_TRY(ensure_io(h, n));
  const int k = h->cur;
  h->cur = (k + 1) % dqnb_sumtree::kRing;
  if (h->busy[k]) { DQNB_CUDA(cudaEventSynchronize(h->ev[k])); h->busy[k] = false; }
  *out = k;
  return DQNB_OK;
}
}  // namespace

extern "C" int dqnb_sumtree_create(int64_t first_leaf, int device, dqnb_sumtree** out) {
  DQNB_REQUIRE(out != nullptr, "out is null");
  DQNB_REQUIRE(first_leaf >= 1 && first_leaf <= ((int64_t)1 << 30) && (first_leaf & (first_leaf - 1)) == 0,
               "first_leaf %lld is not a power of two in [1, 2^30]", (long long)first_leaf);
  DeviceGuard guard(device);
  dqnb_sumtree* h = new dqnb_sumtree();
  h->device = device;
  h->first_leaf = first_leaf;
  while (((int64_t)1 << h->levels) < first_leaf) h->levels++;
  int rc = DQNB_OK;
  if (cudaMalloc(&h->st, sizeof(double) * 2 * (size_t)first_leaf) != cudaSuccess ||
      cudaMemset(h->st, 0, sizeof(double) * 2 * (size_t)first_leaf) != cudaSuccess ||
      cudaMalloc(&h->d_flag, sizeof(int)) != cudaSuccess || cudaMemset(h->d_flag, 0, sizeof(int)) != cudaSuccess) {
    set_error("sum tree of %lld leaves: %s", (long long)first_leaf, cudaGetErrorString(cudaGetLastError()));
    rc = DQNB_ENOMEM;
  }
  for (auto& e : h->ev)
    if (rc == DQNB_OK && cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) rc = DQNB_ECUDA;
  if (rc != DQNB_OK) { dqnb_sumtree_destroy(h); return rc; }
  *out = h;
  return DQNB_OK;
}

extern "C" int dqnb_sumtree_destroy(dqnb_sumtree* h) {
  if (!h) return DQNB_OK;
  DeviceGuard guard(h->device);
  cudaDeviceSynchronize();
  cudaFree(h->st); cudaFree(h->d_flag);
  for (int k = 0; k < dqnb_sumtree::kRing; ++k) {
    if (h->pinned[k]) cudaFreeHost(h->pinned[k]);
    cudaFree(h->dev[k]);
    if (h->ev[k]) cudaEventDestroy(h->ev[k]);
  }
  delete h;
  return DQNB_OK;
}

extern "C" int dqnb_sumtree_set(dqnb_sumtree* h, const int64_t* indices, const double* values, int64_t n, uintptr_t stream) {
  DQNB_REQUIRE(h && (n == 0 || (indices && values)), "null argument");
  if (n == 0) return DQNB_OK;
  for (int64_t i = 0; i < n; ++i) {
    DQNB_REQUIRE(indices[i] >= 0 && indices[i] < h->first_leaf, "leaf %lld outside [0,%lld)", (long long)indices[i],
                 (long long)h->first_leaf);
    DQNB_REQUIRE(values[i] >= 0.0 && values[i] <= 1.7976931348623157e308, "value must be finite and positive.");
  }
  DeviceGuard guard(h->device);
  cudaStream_t st = as_stream(stream);
  int k;
  DQNB_TRY(take_block(h, n, &k));
  int64_t* pi = reinterpret_cast<int64_t*>(h->pinned[k]);
  double* pv = reinterpret_cast<double*>(h->pinned[k] + 8 * (size_t)h->io_cap);
  // duplicate indices: NumPy's `values[indices] = v` keeps the LAST one (replay.py:407); the kernel would race
  int64_t m = 0;
  for (int64_t i = 0; i < n; ++i) {
    bool later = false;
    for (int64_t j = i + 1; j < n && !later; ++j) later = indices[j] == indices[i];
    if (!later) { pi[m] = indices[i]; pv[m] = values[i]; ++m; }
  }
  DQNB_CUDA(cudaMemcpyAsync(h->dev[k], h->pinned[k], 16 * (size_t)h->io_cap, cudaMemcpyHostToDevice, st));
  sumtree_set_kernel<<<1, kSetThreads, 0, st>>>(h->st, h->first_leaf, h->levels,
                                               reinterpret_cast<const int64_t*>(h->dev[k]),
                                               reinterpret_cast<const double*>(h->dev[k] + 8 * (size_t)h->io_cap), (int)m);
  DQNB_CUDA(cudaGetLastError());
  DQNB_CUDA(cudaEventRecord(h->ev[k], st));
  h->busy[k] = true;
  return DQNB_OK;
}

extern "C" int dqnb_sumtree_query(dqnb_sumtree* h, const double* targets_or_uniforms, int64_t n, int scale_by_root,
                                  int64_t* indices, double* root, uintptr_t stream) {
  DQNB_REQUIRE(h && (n == 0 || (targets_or_uniforms && indices)), "null argument");
  DeviceGuard guard(h->device);
  cudaStream_t st = as_stream(stream);
  int k;
  DQNB_TRY(take_block(h, std::max<int64_t>(n, 1), &k));
  double* pu = reinterpret_cast<double*>(h->pinned[k]);
  int64_t* po = reinterpret_cast<int64_t*>(h->pinned[k] + 8 * (size_t)h->io_cap);
  if (n) memcpy(pu, targets_or_uniforms, sizeof(double) * n);
  double* du = reinterpret_cast<double*>(h->dev[k]);
  int64_t* dout = reinterpret_cast<int64_t*>(h->dev[k] + 8 * (size_t)h->io_cap);
  double* droot = reinterpret_cast<double*>(h->dev[k] + 16 * (size_t)h->io_cap);
  if (n) DQNB_CUDA(cudaMemcpyAsync(du, pu, sizeof(double) * n, cudaMemcpyHostToDevice, st));
  sumtree_query_kernel<<<(unsigned)((std::max<int64_t>(n, 1) + 127) / 128), 128, 0, st>>>(
      h->st, h->first_leaf, du, (int)n, scale_by_root, dout, droot, h->d_flag);
  DQNB_CUDA(cudaGetLastError());
  // indices, root and the range flag come back in one block
  if (n) DQNB_CUDA(cudaMemcpyAsync(po, dout, sizeof(int64_t) * n, cudaMemcpyDeviceToHost, st));
  double* proot = reinterpret_cast<double*>(h->pinned[k] + 16 * (size_t)h->io_cap);
  DQNB_CUDA(cudaMemcpyAsync(proot, droot, sizeof(double), cudaMemcpyDeviceToHost, st));
  int* pflag = reinterpret_cast<int*>(h->pinned[k] + 16 * (size_t)h->io_cap + 8);
  DQNB_CUDA(cudaMemcpyAsync(pflag, h->d_flag, sizeof(int), cudaMemcpyDeviceToHost, st));
  DQNB_CUDA(cudaStreamSynchronize(st));
  if (*pflag) {
    DQNB_CUDA(cudaMemsetAsync(h->d_flag, 0, sizeof(int), st));
    set_error("Require 0 <= target < total sum.");
    return DQNB_EINVAL;
  }
  if (n) memcpy(indices, po, sizeof(int64_t) * n);
  if (root) *root = *proot;
  return DQNB_OK;
}

// whole-array copies: storage[2 * first_leaf] doubles (get_state / set_state, resize, set_all)
extern "C" int dqnb_sumtree_read(dqnb_sumtree* h, double* storage_host, int64_t n) {
  DQNB_REQUIRE(h && storage_host && n == 2 * h->first_leaf, "bad argument");
  DeviceGuard guard(h->device);
  DQNB_CUDA(cudaMemcpy(storage_host, h->st, sizeof(double) * n, cudaMemcpyDeviceToHost));
  return DQNB_OK;
}
extern "C" int dqnb_sumtree_write(dqnb_sumtree* h, const double* storage_host, int64_t n) {
  DQNB_REQUIRE(h && storage_host && n == 2 * h->first_leaf, "bad argument");
  DeviceGuard guard(h->device);
  DQNB_CUDA(cudaMemcpy(h->st, storage_host, sizeof(double) * n, cudaMemcpyHostToDevice));
  return DQNB_OK;
}
