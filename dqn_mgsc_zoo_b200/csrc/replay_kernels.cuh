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

// Warp 0 of a gather CTA (every lane computes the same, warp-uniform values; `leader` = the elect.sync lane, which
// alone touches shared memory and issues - a `threadIdx.x == 0` test made ptxas wrap every bulk copy in a
// R2UR.BROADCAST waterfall loop): dedupe N plane references (registers, fully unrolled) into buffers 0.., one bulk
// copy per unique plane completing on `bar`; buf[k] = buffer of reference k, -1 = the zero plane.
template <int N>
__device__ __forceinline__ void gather_issue(const int (&ids)[N], int* buf, uint8_t* smem, int chunk, int len,
                                             const uint8_t* planes, int64_t plane_stride, int px0, uint64_t* bar,
                                             bool leader) {
  int uniq[N];
  int nu = 0;
#pragma unroll
  for (int k = 0; k < N; ++k) {
    int b = -1;
    if (ids[k] >= 0) {
#pragma unroll
      for (int j = 0; j < N; ++j)
        if (j < nu && uniq[j] == ids[k]) b = j;
      if (b < 0) {
        b = nu;
#pragma unroll
        for (int j = 0; j < N; ++j)
          if (j == nu) uniq[j] = ids[k];
        ++nu;
      }
    }
    if (leader) buf[k] = b;
  }
  if (leader) {
    if (nu > 0) {
      mbar_expect_tx(bar, (uint32_t)(nu * len));
#pragma unroll
      for (int j = 0; j < N; ++j)
        if (j < nu) bulk_g2s(smem + j * chunk, planes + (int64_t)uniq[j] * plane_stride + px0, (uint32_t)len, bar);
    } else {
      mbar_arrive(bar);
    }
  }
}

// `cap_planes` = plane buffers the launch gave each CTA (8, or 5 in the throughput regime: a chained transition has
// 5 unique planes, and 5 x chunk of shared memory lets 6 CTAs share an SM instead of 4 - more loads in flight).  A row
// with more unique planes than that (reservoir slots, episode joints) runs in two phases: the s_tm1 stack, then s_t.
template <int kThreads>
__global__ void __launch_bounds__(kThreads)
gather_stack4(ReplayDeviceView v, const int32_t* __restrict__ slots, const __grid_constant__ InlineSlots inl, int n,
              int chunk, int cap_planes, uint8_t* __restrict__ s_tm1_out, uint8_t* __restrict__ s_t_out,
              int32_t* __restrict__ a_out, float* __restrict__ r_out, float* __restrict__ d_out) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ int buf_of[8];   // smem buffer index per plane reference, -1 = zeros
  __shared__ int late_ids[4]; // two-phase rows: the s_t plane references, issued after the s_tm1 stack is out
  __shared__ int two_phase;
  const int row = blockIdx.y, split = blockIdx.x;
  const int slot = slots != nullptr ? slots[row] : inl.v[row];
  const int px0 = split * chunk;
  const int len = min(chunk, v.plane_bytes - px0);   // multiple of 16 by construction

  if (threadIdx.x < 32) {
    const bool leader = elect_one();
    if (leader) {
      mbar_init(&bar, 1);
      fence_mbar_init();
    }
    __syncwarp();
    const int4* idp = reinterpret_cast<const int4*>(v.slot_ids + (int64_t)slot * 8);
    int4 lo = idp[0], hi = idp[1];
    const int ids[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
    int nu = 8;
    if (cap_planes < 8) {      // unique non-zero references (only when the launch is short of buffers)
      nu = 0;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        bool fresh = ids[k] >= 0;
#pragma unroll
        for (int j = 0; j < k; ++j) fresh = fresh && ids[j] != ids[k];
        nu += fresh ? 1 : 0;
      }
    }
    const bool split_row = cap_planes < 8 && nu > cap_planes;
    if (leader) two_phase = split_row;
    if (!split_row) {
      gather_issue<8>(ids, buf_of, smem, chunk, len, v.planes, v.plane_stride, px0, &bar, leader);
    } else {
      const int first[4] = {ids[0], ids[1], ids[2], ids[3]};
      gather_issue<4>(first, buf_of, smem, chunk, len, v.planes, v.plane_stride, px0, &bar, leader);
      if (leader) {
#pragma unroll
        for (int k = 0; k < 4; ++k) late_ids[k] = ids[4 + k];
      }
    }
    if (split == 0 && leader) {
      a_out[row] = v.actions[slot];
      r_out[row] = v.rewards[slot];
      d_out[row] = v.discounts[slot];
    }
  }
  __syncthreads();
  mbar_wait(&bar, 0);

  const int64_t out_base = ((int64_t)row * v.plane_bytes + px0) * 4;
  uint4* o1 = reinterpret_cast<uint4*>(s_tm1_out + out_base);
  uint4* o2 = reinterpret_cast<uint4*>(s_t_out + out_base);
  const int nq = len >> 2;   // groups of 4 pixels
  if (!two_phase) {
    const uint32_t* w[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) w[k] = buf_of[k] >= 0 ? reinterpret_cast<const uint32_t*>(smem + buf_of[k] * chunk) : nullptr;
    for (int q = threadIdx.x; q < nq; q += kThreads) {
      uint32_t x[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) x[k] = w[k] ? w[k][q] : 0u;
      __stcs(o1 + q, interleave4(x[0], x[1], x[2], x[3]));
      __stcs(o2 + q, interleave4(x[4], x[5], x[6], x[7]));
    }
    return;
  }
#pragma unroll 1
  for (int phase = 0; phase < 2; ++phase) {
    if (phase == 1) {
      __syncthreads();                       // every thread has read the s_tm1 planes
      if (threadIdx.x < 32) {
        const int second[4] = {late_ids[0], late_ids[1], late_ids[2], late_ids[3]};
        gather_issue<4>(second, buf_of + 4, smem, chunk, len, v.planes, v.plane_stride, px0, &bar, elect_one());
      }
      __syncthreads();
      mbar_wait(&bar, 1);
    }
    const uint32_t* w[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int b = buf_of[4 * phase + k];
      w[k] = b >= 0 ? reinterpret_cast<const uint32_t*>(smem + b * chunk) : nullptr;
    }
    uint4* o = phase ? o2 : o1;
    for (int q = threadIdx.x; q < nq; q += kThreads) {
      uint32_t x[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) x[k] = w[k] ? w[k][q] : 0u;
      __stcs(o + q, interleave4(x[0], x[1], x[2], x[3]));
    }
  }
}

// Large launches (thousands of transitions: the data-parallel learner's batch, prefetch of many minibatches): the
// one-CTA-per-transition kernel above spends most of a CTA's life in three DEPENDENT memory round trips (slot ->
// plane ids -> planes) before its store phase, with only 4 CTAs per SM to hide them (56 KB of smem each).  Here one
// persistent CTA per SM walks transitions row = blockIdx.x, + gridDim.x, ... through a kGatherStages-deep ring of
// whole-transition buffers: while the CTA stores transition k, the bulk copies of k+1 are in flight and the plane
// ids of k+2 are being fetched, so every round trip is hidden behind a store phase.
constexpr int kGatherStages = 3;
constexpr int kGatherConsumers = 512;                       // warps 0-15 store; warp 16 lane 0 produces
constexpr int kGatherPipeThreads = kGatherConsumers + 32;
constexpr int kGatherMaxItems = 1024;                       // transitions per CTA (slot list staged in smem)
__global__ void __launch_bounds__(kGatherPipeThreads, 1)
gather_stack4_pipe(ReplayDeviceView v, const int32_t* __restrict__ slots, int n, uint8_t* __restrict__ s_tm1_out,
                   uint8_t* __restrict__ s_t_out, int32_t* __restrict__ a_out, float* __restrict__ r_out,
                   float* __restrict__ d_out) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t full[kGatherStages], empty[kGatherStages];
  __shared__ int buf_of[kGatherStages][8];
  __shared__ int slot_s[kGatherMaxItems];
  const int plane = v.plane_bytes;                  // multiple of 16
  const int stage_bytes = 8 * plane;
  const int t = threadIdx.x;
  const int first = blockIdx.x, step = gridDim.x;
  const int my = first < n ? (n - first + step - 1) / step : 0;     // transitions of this CTA (<= kGatherMaxItems)
  if (t == 0) {
    for (int s = 0; s < kGatherStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    fence_mbar_init();
  }
  for (int i = t; i < my; i += kGatherPipeThreads) slot_s[i] = slots[first + i * step];
  __syncthreads();

  if (t >= kGatherConsumers) {
    // ------------------------------------------------------------------ producer (one lane)
    if (t != kGatherConsumers) return;
    // two register sets: everything item j+2 needs is requested while item j is being issued, so the only
    // memory round trip the producer ever waits for was started two items earlier
    int4 lo[2], hi[2];
    int32_t act[2];
    float rew[2], dis[2];
    auto fetch = [&](int i, int set) {
      if (i >= my) return;
      const int slot = slot_s[i];
      const int4* idp = reinterpret_cast<const int4*>(v.slot_ids + (int64_t)slot * 8);
      lo[set] = __ldg(idp); hi[set] = __ldg(idp + 1);
      act[set] = __ldg(v.actions + slot); rew[set] = __ldg(v.rewards + slot); dis[set] = __ldg(v.discounts + slot);
    };
    auto issue = [&](int i, int set) {
      const int stage = i % kGatherStages;
      if (i >= kGatherStages) mbar_wait(&empty[stage], (uint32_t)((i / kGatherStages) - 1) & 1u);
      const int row = first + i * step;
      const int ids[8] = {lo[set].x, lo[set].y, lo[set].z, lo[set].w, hi[set].x, hi[set].y, hi[set].z, hi[set].w};
      int uniq[8];
      int nu = 0;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        int b = -1;
        if (ids[k] >= 0) {
          for (int q = 0; q < nu; ++q)
            if (uniq[q] == ids[k]) b = q;
          if (b < 0) { b = nu; uniq[nu++] = ids[k]; }
        }
        buf_of[stage][k] = b;
      }
      a_out[row] = act[set]; r_out[row] = rew[set]; d_out[row] = dis[set];
      if (nu > 0) {
        mbar_expect_tx(&full[stage], (uint32_t)(nu * plane));
        for (int q = 0; q < nu; ++q)
          bulk_g2s(smem + stage * stage_bytes + q * plane, v.planes + (int64_t)uniq[q] * v.plane_stride,
                   (uint32_t)plane, &full[stage]);
      } else {
        mbar_arrive(&full[stage]);
      }
    };
    fetch(0, 0);
    fetch(1, 1);
    for (int i = 0; i < my; i += 2) {
      issue(i, 0);
      fetch(i + 2, 0);
      if (i + 1 < my) {
        issue(i + 1, 1);
        fetch(i + 3, 1);
      }
    }
    return;
  }
  // -------------------------------------------------------------------- consumers
  for (int k = 0; k < my; ++k) {
    const int stage = k % kGatherStages;
    const int row = first + k * step;
    mbar_wait(&full[stage], (uint32_t)(k / kGatherStages) & 1u);
    const uint8_t* base = smem + stage * stage_bytes;
    const uint32_t* w[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) w[j] = buf_of[stage][j] >= 0 ? reinterpret_cast<const uint32_t*>(base + buf_of[stage][j] * plane) : nullptr;
    const int64_t out_base = (int64_t)row * plane * 4;
    uint4* o1 = reinterpret_cast<uint4*>(s_tm1_out + out_base);
    uint4* o2 = reinterpret_cast<uint4*>(s_t_out + out_base);
    const int nq = plane >> 2;
#pragma unroll 2
    for (int q = t; q < nq; q += kGatherConsumers) {
      uint32_t x[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) x[j] = w[j] ? w[j][q] : 0u;
      __stcs(o1 + q, interleave4(x[0], x[1], x[2], x[3]));
      __stcs(o2 + q, interleave4(x[4], x[5], x[6], x[7]));
    }
    asm volatile("bar.sync 1, %0;" ::"n"(kGatherConsumers) : "memory");   // every consumer has read the stage
    if (t == 0) mbar_arrive(&empty[stage]);
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
