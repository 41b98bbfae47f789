// K1: replay frame-store kernels (HBM-bound byte work; no tensor cores).
//   commit_kernel : scatter staged planes + slot records into the pool / slot tables
//   gather_stack4 : gather + 4-frame stack via 1-D bulk-async (TMA engine) copies into smem,
//                   4x4 byte transpose with PRMT, 16-byte coalesced HWC stores
//   gather_generic: any plane size / stack depth (small test shapes)
#pragma once
#include "common.cuh"

namespace dqnb {

struct SlotRecord {      // one per staged put, 64 bytes
  int32_t slot;
  int32_t action;
  float reward;
  float discount;
  int32_t plane_ids[8];  // s_tm1 planes [0..S), s_t planes [S..2S); -1 = zero plane
  int32_t pad[4];
};

struct ReplayDeviceView {
  const uint8_t* planes;     // [pool][plane_stride]
  const int32_t* slot_ids;   // [capacity][8]
  const int32_t* actions;    // [capacity]
  const float* rewards;      // [capacity]
  const float* discounts;    // [capacity]
  int64_t plane_stride;      // bytes between planes (multiple of 16)
  int plane_bytes;           // H*W
  int stack;                 // S
};

// Staged uploads -> pool.  grid.x = n_planes + ceil(n_records/kCommitThreads).  The staged data
// may sit in pinned host memory (read over PCIe), so a thread issues all its loads before the
// first store.
constexpr int kCommitThreads = 256;
__global__ void __launch_bounds__(kCommitThreads) commit_kernel(uint8_t* __restrict__ planes, int64_t plane_stride, int plane_bytes,
                              const uint8_t* __restrict__ staged_planes, const int32_t* __restrict__ staged_dst,
                              int n_planes, const SlotRecord* __restrict__ records, int n_records,
                              int32_t* __restrict__ slot_ids, int32_t* __restrict__ actions,
                              float* __restrict__ rewards, float* __restrict__ discounts) {
  int b = blockIdx.x;
  if (b < n_planes) {
    const uint4* src = reinterpret_cast<const uint4*>(staged_planes + (int64_t)b * plane_stride);
    uint4* dst = reinterpret_cast<uint4*>(planes + (int64_t)staged_dst[b] * plane_stride);
    int n16 = (int)(plane_stride >> 4);
    for (int i0 = threadIdx.x; i0 < n16; i0 += 4 * kCommitThreads) {
      uint4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (i0 + u * kCommitThreads < n16) v[u] = src[i0 + u * kCommitThreads];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (i0 + u * kCommitThreads < n16) dst[i0 + u * kCommitThreads] = v[u];
    }
    return;
  }
  int r = (b - n_planes) * blockDim.x + threadIdx.x;
  if (r < n_records) {
    SlotRecord rec = records[r];
    int4* ids = reinterpret_cast<int4*>(slot_ids + (int64_t)rec.slot * 8);
    ids[0] = make_int4(rec.plane_ids[0], rec.plane_ids[1], rec.plane_ids[2], rec.plane_ids[3]);
    ids[1] = make_int4(rec.plane_ids[4], rec.plane_ids[5], rec.plane_ids[6], rec.plane_ids[7]);
    actions[rec.slot] = rec.action;
    rewards[rec.slot] = rec.reward;
    discounts[rec.slot] = rec.discount;
  }
}

// 4 words holding 4 pixels of channels 0..3  ->  4 words holding pixels 0..3 in HWC order.
__device__ __forceinline__ uint4 interleave4(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3) {
  uint32_t t0 = __byte_perm(w0, w1, 0x5140), t1 = __byte_perm(w2, w3, 0x5140);
  uint32_t t2 = __byte_perm(w0, w1, 0x7362), t3 = __byte_perm(w2, w3, 0x7362);
  return make_uint4(__byte_perm(t0, t1, 0x5410), __byte_perm(t0, t1, 0x7632),
                    __byte_perm(t2, t3, 0x5410), __byte_perm(t2, t3, 0x7632));
}

// One CTA gathers `chunk` pixels (chunk % 16 == 0) of one sampled transition:
// blockIdx.x = split, blockIdx.y = batch row.  Unique planes only are fetched (a chained
// transition shares 3 of its 8 plane references), each with one bulk async copy that
// completes on an mbarrier; threads then emit both stacks with 16-byte stores.
// Small batches (the learner's sample of 32) carry their slot list in the kernel parameters:
// no index upload, i.e. no copy-engine operation between the kernels of the training stream.
constexpr int kInlineSlots = 64;
struct InlineSlots { int32_t v[kInlineSlots]; };

template <int kThreads>
__global__ void __launch_bounds__(kThreads)
gather_stack4(ReplayDeviceView v, const int32_t* __restrict__ slots, const __grid_constant__ InlineSlots inl, int n,
              int chunk, uint8_t* __restrict__ s_tm1_out, uint8_t* __restrict__ s_t_out,
              int32_t* __restrict__ a_out, float* __restrict__ r_out, float* __restrict__ d_out) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ int buf_of[8];   // smem buffer index per plane reference, -1 = zeros
  const int row = blockIdx.y, split = blockIdx.x;
  const int slot = slots != nullptr ? slots[row] : inl.v[row];
  const int px0 = split * chunk;
  const int len = min(chunk, v.plane_bytes - px0);   // multiple of 16 by construction

  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
    const int4* idp = reinterpret_cast<const int4*>(v.slot_ids + (int64_t)slot * 8);
    int4 lo = idp[0], hi = idp[1];
    int ids[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
    int uniq[8];
    int nu = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      int b = -1;
      if (ids[k] >= 0) {
        for (int j = 0; j < nu; ++j)
          if (uniq[j] == ids[k]) b = j;
        if (b < 0) {
          b = nu;
          uniq[nu++] = ids[k];
        }
      }
      buf_of[k] = b;
    }
    if (nu > 0) {
      mbar_expect_tx(&bar, (uint32_t)(nu * len));
      for (int j = 0; j < nu; ++j)
        bulk_g2s(smem + j * chunk, v.planes + (int64_t)uniq[j] * v.plane_stride + px0, (uint32_t)len, &bar);
    } else {
      mbar_arrive(&bar);
    }
    if (split == 0) {
      a_out[row] = v.actions[slot];
      r_out[row] = v.rewards[slot];
      d_out[row] = v.discounts[slot];
    }
  }
  __syncthreads();
  mbar_wait(&bar, 0);

  const uint32_t* w[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) w[k] = buf_of[k] >= 0 ? reinterpret_cast<const uint32_t*>(smem + buf_of[k] * chunk) : nullptr;
  const int64_t out_base = ((int64_t)row * v.plane_bytes + px0) * 4;
  uint4* o1 = reinterpret_cast<uint4*>(s_tm1_out + out_base);
  uint4* o2 = reinterpret_cast<uint4*>(s_t_out + out_base);
  const int nq = len >> 2;   // groups of 4 pixels
  for (int q = threadIdx.x; q < nq; q += kThreads) {
    uint32_t x[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) x[k] = w[k] ? w[k][q] : 0u;
    __stcs(o1 + q, interleave4(x[0], x[1], x[2], x[3]));
    __stcs(o2 + q, interleave4(x[4], x[5], x[6], x[7]));
  }
}

// Fallback for shapes the fast path does not cover (stack != 4 or H*W % 16 != 0).
__global__ void gather_generic(ReplayDeviceView v, const int32_t* __restrict__ slots, int n,
                               uint8_t* __restrict__ s_tm1_out, uint8_t* __restrict__ s_t_out,
                               int32_t* __restrict__ a_out, float* __restrict__ r_out, float* __restrict__ d_out) {
  const int row = blockIdx.x;
  const int slot = slots[row];
  const int S = v.stack;
  const int32_t* ids = v.slot_ids + (int64_t)slot * 8;
  const int total = v.plane_bytes * S;
  for (int e = threadIdx.x; e < total; e += blockDim.x) {
    int px = e / S, c = e - px * S;
    int p1 = ids[c], p2 = ids[S + c];
    s_tm1_out[(int64_t)row * total + e] = p1 >= 0 ? v.planes[(int64_t)p1 * v.plane_stride + px] : 0;
    s_t_out[(int64_t)row * total + e] = p2 >= 0 ? v.planes[(int64_t)p2 * v.plane_stride + px] : 0;
  }
  if (threadIdx.x == 0) {
    a_out[row] = v.actions[slot];
    r_out[row] = v.rewards[slot];
    d_out[row] = v.discounts[slot];
  }
}

}  // namespace dqnb
