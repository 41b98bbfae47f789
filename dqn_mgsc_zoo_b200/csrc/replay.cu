// Replay frame store: host bookkeeping (plane pool, dedup, staging) + kernel launches.
// Replaces the item storage of the reference replays (replay.py:163-333,
// replay_circular.py:90-146,568-664,1217-1312); see include/dqnb200.h.
#include <stdlib.h>

#include <algorithm>
#include <vector>

#include "replay_kernels.cuh"

using namespace dqnb;

namespace {
constexpr int kStageRecords = 64;            // puts per staging batch
constexpr int kStagePlanes = 256;            // uploaded planes per staging batch
constexpr int kIdxRing = 8;
constexpr int kGatherThreads = 256;
constexpr size_t kZeroCopyBytes = 128 * 1024;   // staged batches up to this size are read in place by the GPU
constexpr int64_t kZeroCopySlots = 0;           // gather index lists: in-place reads by the gather CTAs measured slower (+7 us at n=32) than the copy

struct Stage {
  uint8_t* host = nullptr;   // pinned: [records][dst ids][planes...]
  uint8_t* dev = nullptr;
  cudaEvent_t copied = nullptr;
  bool in_flight = false;
  int n_planes = 0, n_records = 0;
};
}  // namespace

struct dqnb_replay {
  int device = 0;
  int64_t capacity = 0;
  int H = 0, W = 0, S = 0, plane_bytes = 0;
  int64_t plane_stride = 0;
  // device state
  uint8_t* d_planes = nullptr;
  int64_t pool_planes = 0;
  int32_t* d_slot_ids = nullptr;
  int32_t* d_actions = nullptr;
  float* d_rewards = nullptr;
  float* d_discounts = nullptr;
  // staging
  Stage stage[2];
  int cur = 0;
  size_t hdr_records = 0, hdr_dst = 0, hdr_bytes = 0;
  // host bookkeeping
  std::vector<int32_t> refcount, free_list, pending_free, slot_ids;
  std::vector<uint8_t> slot_staged;      // slot already has a record in the open batch
  std::vector<uint8_t> last_st;
  std::vector<uint8_t> newest;   // scratch: channel 3 of the stack being put, as a plane
  int32_t last_ids[8];
  bool have_last = false;
  // gather index ring + host-return buffers
  int32_t* h_idx = nullptr;
  int32_t* d_idx = nullptr;
  int64_t idx_cap = 0;
  cudaEvent_t idx_ev[kIdxRing] = {};
  bool idx_busy[kIdxRing] = {};
  int idx_cur = 0;
  int idx_pending = -1;                  // ring entry handed to the next gather without a copy
  // small commits run on a side stream so they overlap whatever follows the last gather on the
  // caller's stream (the learner's update graph): ordered after that gather, before the next one
  cudaStream_t side = nullptr;
  cudaEvent_t gather_done = nullptr;
  // callers may gather on different streams (the learner prefetches on its own gather stream): a
  // gather on another stream than its predecessor / than the last commit's waits for them first
  cudaStream_t gather_stream = nullptr, commit_stream = nullptr;
  cudaStream_t commit_ran_on = nullptr;   // the stream the last commit kernel really ran on (side or caller's)
  cudaEvent_t committed = nullptr;
  bool have_commit = false;
  bool have_gather = false;
  uint8_t* d_out = nullptr;
  int64_t out_cap = 0;
  int gather_splits = 0;
  int64_t n_puts = 0, n_uploaded = 0, n_chained = 0, n_growths = 0;

  SlotRecord* rec_of(Stage& s) { return reinterpret_cast<SlotRecord*>(s.host); }
  int32_t* dst_of(Stage& s) { return reinterpret_cast<int32_t*>(s.host + hdr_records); }
  uint8_t* planes_of(Stage& s) { return s.host + hdr_bytes; }
};

namespace {

int grow_pool(dqnb_replay* h, int64_t want) {
  // Rare: reallocate the pool and copy (device-wide sync).  Plane ids stay valid.
  int64_t new_planes = std::max<int64_t>(want, h->pool_planes + h->pool_planes / 2 + 64);
  uint8_t* np = nullptr;
  cudaError_t e = cudaMalloc(&np, (size_t)new_planes * h->plane_stride);
  if (e != cudaSuccess) {
    set_error("plane pool growth to %lld planes failed: %s", (long long)new_planes, cudaGetErrorString(e));
    return DQNB_ENOMEM;
  }
  DQNB_CUDA(cudaDeviceSynchronize());
  if (h->d_planes) {
    DQNB_CUDA(cudaMemcpy(np, h->d_planes, (size_t)h->pool_planes * h->plane_stride, cudaMemcpyDeviceToDevice));
    DQNB_CUDA(cudaFree(h->d_planes));
  }
  h->d_planes = np;
  h->refcount.resize(new_planes, 0);
  for (int64_t p = new_planes - 1; p >= h->pool_planes; --p) h->free_list.push_back((int32_t)p);
  h->pool_planes = new_planes;
  h->n_growths++;
  return DQNB_OK;
}

int flush(dqnb_replay* h, cudaStream_t st) {
  Stage& s = h->stage[h->cur];
  if (s.n_records == 0 && s.n_planes == 0) return DQNB_OK;
  size_t bytes = h->hdr_bytes + (size_t)s.n_planes * h->plane_stride;
  // A handful of planes (the steady state: ~1 new frame per put) is read by the commit kernel
  // straight from the pinned staging buffer over PCIe; a copy-engine hop between two kernels of
  // the training stream costs more than the transfer.  Big batches (initial fill) are copied.
  const bool zero_copy = bytes <= kZeroCopyBytes;
  const uint8_t* src = zero_copy ? s.host : s.dev;
  if (!zero_copy) {
    DQNB_CUDA(cudaMemcpyAsync(s.dev, s.host, bytes, cudaMemcpyHostToDevice, st));
    DQNB_CUDA(cudaEventRecord(s.copied, st));
  }
  int grid = s.n_planes + (s.n_records + kCommitThreads - 1) / kCommitThreads;
  // Readers of what this commit overwrites (evicted slot records, recycled planes) are gathers
  // already enqueued on `st`; everything enqueued after the last of them (the update graph) does
  // not touch the replay, so a small commit may run beside it.
  cudaStream_t cs = (zero_copy && h->side != nullptr) ? h->side : st;
  if (h->have_gather && (cs != st || st != h->gather_stream)) DQNB_CUDA(cudaStreamWaitEvent(cs, h->gather_done, 0));
  // Commits must land in issue order (two batches may write the same slot: small capacities, reservoir
  // replacements): a commit on another stream than its predecessor waits for it.
  if (h->have_commit && cs != h->commit_ran_on) DQNB_CUDA(cudaStreamWaitEvent(cs, h->committed, 0));
  commit_kernel<<<grid, kCommitThreads, 0, cs>>>(
      h->d_planes, h->plane_stride, h->plane_bytes, src + h->hdr_bytes,
      reinterpret_cast<const int32_t*>(src + h->hdr_records), s.n_planes,
      reinterpret_cast<const SlotRecord*>(src), s.n_records, h->d_slot_ids, h->d_actions,
      h->d_rewards, h->d_discounts);
  DQNB_CUDA(cudaGetLastError());
  if (zero_copy) {
    DQNB_CUDA(cudaEventRecord(s.copied, cs));   // staging is free once the kernel has read it
    if (cs != st) DQNB_CUDA(cudaStreamWaitEvent(st, s.copied, 0));   // later work on st sees the commit
  }
  DQNB_CUDA(cudaEventRecord(h->committed, cs));   // gathers issued on any other stream wait for this
  h->commit_stream = st;
  h->commit_ran_on = cs;
  h->have_commit = true;
  s.in_flight = true;
  for (int i = 0; i < s.n_records; ++i) h->slot_staged[h->rec_of(s)[i].slot] = 0;
  // planes released during this batch may be handed out again only now (stream order
  // puts their next write after every earlier reader).
  h->free_list.insert(h->free_list.end(), h->pending_free.begin(), h->pending_free.end());
  h->pending_free.clear();
  s.n_planes = s.n_records = 0;
  h->cur ^= 1;
  Stage& nx = h->stage[h->cur];
  if (nx.in_flight) {
    DQNB_CUDA(cudaEventSynchronize(nx.copied));
    nx.in_flight = false;
  }
  return DQNB_OK;
}

inline void release_plane(dqnb_replay* h, int32_t p) {
  if (p >= 0 && --h->refcount[p] == 0) h->pending_free.push_back(p);
}

// Copies channel c of an HWC stack into a staged plane; returns the new plane id.
// `extracted`: channel c of `stack` already pulled out as a plane (match_copy_extract), or null.
int stage_plane(dqnb_replay* h, const uint8_t* stack, int c, cudaStream_t st, int32_t* out_id,
                const uint8_t* extracted = nullptr) {
  if (h->stage[h->cur].n_planes == kStagePlanes) DQNB_TRY(flush(h, st));
  if (h->free_list.empty()) {
    DQNB_TRY(flush(h, st));   // recycles pending_free
    if (h->free_list.empty()) DQNB_TRY(grow_pool(h, 0));
  }
  int32_t p = h->free_list.back();
  h->free_list.pop_back();
  h->refcount[p] = 1;
  Stage& s = h->stage[h->cur];
  uint8_t* dst = h->planes_of(s) + (size_t)s.n_planes * h->plane_stride;
  const int n = h->plane_bytes, S = h->S;
  if (extracted != nullptr) {
    memcpy(dst, extracted, (size_t)n);
  } else if (S == 4) {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(stack);
    const int sh = 8 * c;
    for (int i = 0; i < n; ++i) dst[i] = (uint8_t)(src[i] >> sh);
  } else {
    for (int i = 0; i < n; ++i) dst[i] = stack[(size_t)i * S + c];
  }
  h->dst_of(s)[s.n_planes] = p;
  s.n_planes++;
  h->n_uploaded++;
  *out_id = p;
  return DQNB_OK;
}

// For each channel of `cur`: -1 if all zero, d>=0 if equal to channel d of `prev`, -2 if new.
void match_channels(const uint8_t* cur, const uint8_t* prev, int npx, int S, int* match) {
  if (S == 4) {
    const uint32_t* a = reinterpret_cast<const uint32_t*>(cur);
    const uint32_t* b = prev ? reinterpret_cast<const uint32_t*>(prev) : nullptr;
    uint32_t any = 0, dshift = 0, dsame = 0;
    if (b) {
      for (int i = 0; i < npx; ++i) {
        uint32_t x = a[i], y = b[i];
        any |= x;
        dshift |= (x & 0x00FFFFFFu) ^ (y >> 8);
        dsame |= x ^ y;
      }
    } else {
      for (int i = 0; i < npx; ++i) any |= a[i];
      dshift = dsame = 0xFFFFFFFFu;
    }
    for (int c = 0; c < 4; ++c) {
      if (((any >> (8 * c)) & 0xFF) == 0) match[c] = -1;
      else if (c < 3 && ((dshift >> (8 * c)) & 0xFF) == 0) match[c] = c + 1;
      else if (((dsame >> (8 * c)) & 0xFF) == 0) match[c] = c;
      else match[c] = -2;
    }
    return;
  }
  for (int c = 0; c < S; ++c) {
    bool zero = true;
    for (int i = 0; i < npx && zero; ++i) zero = cur[(size_t)i * S + c] == 0;
    match[c] = zero ? -1 : -2;
    if (zero || !prev) continue;
    for (int d = 0; d < S && match[c] == -2; ++d) {
      bool eq = true;
      for (int i = 0; i < npx && eq; ++i) eq = cur[(size_t)i * S + c] == prev[(size_t)i * S + d];
      if (eq) match[c] = d;
    }
  }
}

// match_channels(cur, prev) for 4-deep stacks in ONE pass that also leaves a copy of `cur` in `copy` (the next put's
// chain check) and the newest channel (3: the frame a chained observation adds) as a plane in `newest`: a put used
// to walk the 28 KB stack four times (compare, match, extract, copy), which was half of its host time.
void match_copy_extract(const uint8_t* cur, const uint8_t* prev, int npx, uint8_t* copy, uint8_t* newest, int* match) {
  const uint32_t* a = reinterpret_cast<const uint32_t*>(cur);
  const uint32_t* b = reinterpret_cast<const uint32_t*>(prev);
  uint32_t* c = reinterpret_cast<uint32_t*>(copy);
  uint32_t any = 0, dshift = 0, dsame = 0;
  for (int i = 0; i < npx; ++i) {
    const uint32_t x = a[i], y = b[i];
    any |= x;
    dshift |= (x & 0x00FFFFFFu) ^ (y >> 8);
    dsame |= x ^ y;
    c[i] = x;
    newest[i] = (uint8_t)(x >> 24);
  }
  for (int ch = 0; ch < 4; ++ch) {
    if (((any >> (8 * ch)) & 0xFF) == 0) match[ch] = -1;
    else if (ch < 3 && ((dshift >> (8 * ch)) & 0xFF) == 0) match[ch] = ch + 1;
    else if (((dsame >> (8 * ch)) & 0xFF) == 0) match[ch] = ch;
    else match[ch] = -2;
  }
}

}  // namespace

// ------------------------------------------------------------------------- C ABI
extern "C" int dqnb_replay_create(int64_t capacity, int height, int width, int stack,
                                  int64_t initial_planes, int device, dqnb_replay** out) {
  DQNB_REQUIRE(out != nullptr, "out is null");
  DQNB_REQUIRE(capacity > 0 && capacity < (1LL << 31), "capacity %lld out of range", (long long)capacity);
  DQNB_REQUIRE(height > 0 && width > 0, "bad frame shape %dx%d", height, width);
  DQNB_REQUIRE(stack >= 1 && stack <= 4, "stack depth %d not in 1..4", stack);
  DeviceGuard guard(device);
  dqnb_replay* h = new dqnb_replay();
  // everything below may fail half way: the handle is destroyed (destroy tolerates partial initialisation)
  const int rc_create = [&]() -> int {
  h->device = device;
  h->capacity = capacity;
  h->H = height; h->W = width; h->S = stack;
  h->plane_bytes = height * width;
  h->plane_stride = (h->plane_bytes + 15) / 16 * 16;
  h->hdr_records = sizeof(SlotRecord) * kStageRecords;
  h->hdr_dst = sizeof(int32_t) * kStagePlanes;
  h->hdr_bytes = h->hdr_records + h->hdr_dst;
  h->slot_ids.assign((size_t)capacity * 8, -1);
  h->slot_staged.assign((size_t)capacity, 0);
  h->last_st.resize((size_t)h->plane_bytes * stack);
  DQNB_CUDA(cudaMalloc(&h->d_slot_ids, sizeof(int32_t) * 8 * capacity));
  DQNB_CUDA(cudaMemset(h->d_slot_ids, 0xFF, sizeof(int32_t) * 8 * capacity));
  DQNB_CUDA(cudaMalloc(&h->d_actions, sizeof(int32_t) * capacity));
  DQNB_CUDA(cudaMalloc(&h->d_rewards, sizeof(float) * capacity));
  DQNB_CUDA(cudaMalloc(&h->d_discounts, sizeof(float) * capacity));
  DQNB_CUDA(cudaMemset(h->d_actions, 0, sizeof(int32_t) * capacity));
  DQNB_CUDA(cudaMemset(h->d_rewards, 0, sizeof(float) * capacity));
  DQNB_CUDA(cudaMemset(h->d_discounts, 0, sizeof(float) * capacity));
  size_t stage_bytes = h->hdr_bytes + (size_t)kStagePlanes * h->plane_stride;
  for (Stage& s : h->stage) {
    DQNB_CUDA(cudaMallocHost(&s.host, stage_bytes));
    DQNB_CUDA(cudaMalloc(&s.dev, stage_bytes));
    DQNB_CUDA(cudaEventCreateWithFlags(&s.copied, cudaEventDisableTiming));
  }
  for (int i = 0; i < kIdxRing; ++i) DQNB_CUDA(cudaEventCreateWithFlags(&h->idx_ev[i], cudaEventDisableTiming));
  DQNB_CUDA(cudaStreamCreateWithFlags(&h->side, cudaStreamNonBlocking));
  DQNB_CUDA(cudaEventCreateWithFlags(&h->gather_done, cudaEventDisableTiming));
  DQNB_CUDA(cudaEventCreateWithFlags(&h->committed, cudaEventDisableTiming));
  if (initial_planes <= 0) initial_planes = capacity + capacity / 8 + 64;
  DQNB_TRY(grow_pool(h, initial_planes));
  h->n_growths = 0;
  DQNB_CUDA(cudaFuncSetAttribute(gather_stack4<kGatherThreads>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 8 * 16 * 1024));
  if ((size_t)kGatherStages * 8 * h->plane_bytes <= 200 * 1024)
    DQNB_CUDA(cudaFuncSetAttribute(gather_stack4_pipe, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   kGatherStages * 8 * h->plane_bytes));
  return DQNB_OK;
  }();
  if (rc_create != DQNB_OK) { dqnb_replay_destroy(h); return rc_create; }
  *out = h;
  return DQNB_OK;
}

extern "C" int dqnb_replay_destroy(dqnb_replay* h) {
  if (!h) return DQNB_OK;
  DeviceGuard guard(h->device);
  cudaDeviceSynchronize();
  cudaFree(h->d_planes); cudaFree(h->d_slot_ids); cudaFree(h->d_actions);
  cudaFree(h->d_rewards); cudaFree(h->d_discounts); cudaFree(h->d_idx); cudaFree(h->d_out);
  if (h->h_idx) cudaFreeHost(h->h_idx);
  for (Stage& s : h->stage) {
    if (s.host) cudaFreeHost(s.host);
    cudaFree(s.dev);
    if (s.copied) cudaEventDestroy(s.copied);
  }
  for (int i = 0; i < kIdxRing; ++i) if (h->idx_ev[i]) cudaEventDestroy(h->idx_ev[i]);
  if (h->side) cudaStreamDestroy(h->side);
  if (h->gather_done) cudaEventDestroy(h->gather_done);
  if (h->committed) cudaEventDestroy(h->committed);
  delete h;
  return DQNB_OK;
}

extern "C" int dqnb_replay_put(dqnb_replay* h, int64_t slot, const uint8_t* s_tm1, const uint8_t* s_t,
                               int32_t action, float reward, float discount, uintptr_t stream) {
  DQNB_REQUIRE(h && s_tm1 && s_t, "null argument");
  DQNB_REQUIRE(slot >= 0 && slot < h->capacity, "slot %lld outside [0,%lld)", (long long)slot, (long long)h->capacity);
  DeviceGuard guard(h->device);
  cudaStream_t st = as_stream(stream);
  const int S = h->S;
  const size_t stack_bytes = (size_t)h->plane_bytes * S;
  if (h->slot_staged[slot] || h->stage[h->cur].n_records == kStageRecords) DQNB_TRY(flush(h, st));

  int32_t ids[8];
  for (int k = 0; k < 8; ++k) ids[k] = -1;
  // s_tm1: chained to the previous put's s_t (TransitionAccumulator hands over the very
  // same observation, replay.py:1184-1190) -> share its planes; otherwise upload it.
  const bool chained = h->have_last && memcmp(s_tm1, h->last_st.data(), stack_bytes) == 0;
  if (chained) {
    for (int c = 0; c < S; ++c) {
      ids[c] = h->last_ids[c];
      if (ids[c] >= 0) h->refcount[ids[c]]++;
    }
    h->n_chained++;
  } else {
    int m[4];
    match_channels(s_tm1, nullptr, h->plane_bytes, S, m);
    for (int c = 0; c < S; ++c)
      if (m[c] != -1) DQNB_TRY(stage_plane(h, s_tm1, c, st, &ids[c]));
  }
  int m[4];
  const bool one_pass = S == 4;   // (last_st may be overwritten now: the chain check above is done)
  const bool had_last = h->have_last;
  if (one_pass) {
    h->have_last = false;         // until last_ids match the new contents again (an error return below leaves it off)
    if (h->newest.size() < (size_t)h->plane_bytes) h->newest.resize((size_t)h->plane_bytes);
    match_copy_extract(s_t, s_tm1, h->plane_bytes, h->last_st.data(), h->newest.data(), m);
  } else {
    match_channels(s_t, s_tm1, h->plane_bytes, S, m);
  }
  for (int c = 0; c < S; ++c) {
    if (m[c] == -1) continue;
    if (m[c] >= 0 && ids[m[c]] >= 0) {
      ids[S + c] = ids[m[c]];
      h->refcount[ids[S + c]]++;
    } else {
      DQNB_TRY(stage_plane(h, s_t, c, st, &ids[S + c], one_pass && c == 3 ? h->newest.data() : nullptr));
    }
  }
  // release what the slot held, then what `last` held; take refs for the new `last`.
  int32_t* old = &h->slot_ids[(size_t)slot * 8];
  for (int k = 0; k < 8; ++k) release_plane(h, old[k]);
  for (int k = 0; k < 8; ++k) old[k] = ids[k];
  if (had_last)
    for (int c = 0; c < S; ++c) release_plane(h, h->last_ids[c]);
  for (int c = 0; c < S; ++c) {
    h->last_ids[c] = ids[S + c];
    if (ids[S + c] >= 0) h->refcount[ids[S + c]]++;
  }
  if (!one_pass) memcpy(h->last_st.data(), s_t, stack_bytes);
  h->have_last = true;

  Stage& s = h->stage[h->cur];
  SlotRecord& r = h->rec_of(s)[s.n_records++];
  r.slot = (int32_t)slot; r.action = action; r.reward = reward; r.discount = discount;
  for (int k = 0; k < 8; ++k) r.plane_ids[k] = ids[k];
  h->slot_staged[slot] = 1;
  h->n_puts++;
  return DQNB_OK;
}

extern "C" int dqnb_replay_flush(dqnb_replay* h, uintptr_t stream) {
  DQNB_REQUIRE(h, "null handle");
  DeviceGuard guard(h->device);
  return flush(h, as_stream(stream));
}

namespace {

ReplayDeviceView view_of(const dqnb_replay* h) {
  ReplayDeviceView v;
  v.planes = h->d_planes; v.slot_ids = h->d_slot_ids; v.actions = h->d_actions;
  v.rewards = h->d_rewards; v.discounts = h->d_discounts;
  v.plane_stride = h->plane_stride; v.plane_bytes = h->plane_bytes; v.stack = h->S;
  return v;
}

int launch_gather(dqnb_replay* h, const int32_t* slots_dev, int64_t n, uint8_t* s_tm1, uint8_t* s_t,
                  int32_t* a, float* r, float* d, cudaStream_t st, const InlineSlots* inl = nullptr) {
  if (n == 0) return DQNB_OK;
  if (h->have_gather && st != h->gather_stream) DQNB_CUDA(cudaStreamWaitEvent(st, h->gather_done, 0));
  if (h->have_commit && st != h->commit_stream) DQNB_CUDA(cudaStreamWaitEvent(st, h->committed, 0));
  h->gather_stream = st;
  ReplayDeviceView v = view_of(h);
  static const int pipe_min = getenv("DQNB_GATHER_PIPE_MIN") ? atoi(getenv("DQNB_GATHER_PIPE_MIN")) : (1 << 30);   // measured slower than 6 CTAs/SM of the kernel below: off
  const size_t pipe_smem = (size_t)kGatherStages * 8 * h->plane_bytes;
  if (h->S == 4 && h->plane_bytes % 16 == 0 && slots_dev != nullptr && n >= pipe_min && h->gather_splits == 0 &&   // (any explicit split setting selects the per-transition kernel)
      pipe_smem <= 200 * 1024 && n <= (int64_t)kNumSMs * kGatherMaxItems) {
    // thousands of transitions: one persistent CTA per SM, loads of the next transitions behind the stores of this one
    gather_stack4_pipe<<<kNumSMs, kGatherPipeThreads, pipe_smem, st>>>(v, slots_dev, (int)n, s_tm1, s_t, a, r, d);
  } else if (h->S == 4 && h->plane_bytes % 16 == 0) {
    // enough CTAs to cover the 148 SMs twice for small batches; whole planes for big ones
    int splits = h->gather_splits;
    const bool force8 = splits >= 100;          // tuning: 100 + s = s splits (0 = automatic) with 8 plane buffers per CTA
    if (force8) splits -= 100;
    if (splits <= 0) splits = (int)std::min<int64_t>(std::max<int64_t>(1, (2 * kNumSMs + n - 1) / n), 16);
    int chunk = ((h->plane_bytes + splits - 1) / splits + 15) / 16 * 16;
    chunk = std::min(chunk, 16 * 1024);
    splits = (h->plane_bytes + chunk - 1) / chunk;
    dim3 grid(splits, (unsigned)n);
    static const InlineSlots no_inline{};
    // throughput regime: 5 plane buffers per CTA (6 CTAs per SM instead of 4); rows with more unique planes take two phases
    static const int cap5_min = getenv("DQNB_GATHER_CAP5_MIN") ? atoi(getenv("DQNB_GATHER_CAP5_MIN")) : 256;
    const int cap_planes = (n >= cap5_min && !force8) ? 5 : 8;
    gather_stack4<kGatherThreads><<<grid, kGatherThreads, cap_planes * (size_t)chunk, st>>>(
        v, slots_dev, inl ? *inl : no_inline, (int)n, chunk, cap_planes, s_tm1, s_t, a, r, d);
  } else {
    gather_generic<<<(unsigned)n, 128, 0, st>>>(v, slots_dev, (int)n, s_tm1, s_t, a, r, d);
  }
  DQNB_CUDA(cudaGetLastError());
  DQNB_CUDA(cudaEventRecord(h->gather_done, st));
  h->have_gather = true;
  if (h->idx_pending >= 0) {
    DQNB_CUDA(cudaEventRecord(h->idx_ev[h->idx_pending], st));
    h->idx_busy[h->idx_pending] = true;
    h->idx_pending = -1;
  }
  return DQNB_OK;
}

// int64 host slots -> int32 device slots through a small pinned ring.
int upload_slots(dqnb_replay* h, const int64_t* slots, int64_t n, cudaStream_t st, const int32_t** out) {
  if (n > h->idx_cap) {
    DQNB_CUDA(cudaDeviceSynchronize());
    if (h->h_idx) cudaFreeHost(h->h_idx);
    cudaFree(h->d_idx);
    h->idx_cap = std::max<int64_t>(n, 1024);
    DQNB_CUDA(cudaMallocHost(&h->h_idx, sizeof(int32_t) * h->idx_cap * kIdxRing));
    DQNB_CUDA(cudaMalloc(&h->d_idx, sizeof(int32_t) * h->idx_cap * kIdxRing));
    for (bool& b : h->idx_busy) b = false;
  }
  int k = h->idx_cur;
  h->idx_cur = (k + 1) % kIdxRing;
  if (h->idx_busy[k]) DQNB_CUDA(cudaEventSynchronize(h->idx_ev[k]));
  int32_t* hp = h->h_idx + (size_t)k * h->idx_cap;
  int32_t* dp = h->d_idx + (size_t)k * h->idx_cap;
  for (int64_t i = 0; i < n; ++i) {
    if (slots[i] < 0 || slots[i] >= h->capacity) {
      set_error("slot %lld outside [0,%lld)", (long long)slots[i], (long long)h->capacity);
      return DQNB_EINVAL;
    }
    hp[i] = (int32_t)slots[i];
  }
  if (n <= kZeroCopySlots) {   // the gather kernel reads the pinned list in place
    h->idx_pending = k;        // launch_gather records idx_ev[k] behind the kernel
    *out = hp;
    return DQNB_OK;
  }
  DQNB_CUDA(cudaMemcpyAsync(dp, hp, sizeof(int32_t) * n, cudaMemcpyHostToDevice, st));
  DQNB_CUDA(cudaEventRecord(h->idx_ev[k], st));
  h->idx_busy[k] = true;
  *out = dp;
  return DQNB_OK;
}

}  // namespace

extern "C" int dqnb_replay_gather_dev(dqnb_replay* h, const int32_t* slots_dev, int64_t n, uint8_t* s_tm1_dev,
                                      uint8_t* s_t_dev, int32_t* a_dev, float* r_dev, float* disc_dev,
                                      uintptr_t stream) {
  DQNB_REQUIRE(h && slots_dev && s_tm1_dev && s_t_dev && a_dev && r_dev && disc_dev, "null argument");
  DQNB_REQUIRE(n >= 0 && n < 65536 * 16, "n=%lld out of range", (long long)n);
  DeviceGuard guard(h->device);
  cudaStream_t st = as_stream(stream);
  DQNB_TRY(flush(h, st));
  return launch_gather(h, slots_dev, n, s_tm1_dev, s_t_dev, a_dev, r_dev, disc_dev, st);
}

extern "C" int dqnb_replay_gather(dqnb_replay* h, const int64_t* slots_host, int64_t n, uint8_t* s_tm1_dev,
                                  uint8_t* s_t_dev, int32_t* a_dev, float* r_dev, float* disc_dev,
                                  uintptr_t stream) {
  DQNB_REQUIRE(h && (slots_host || n == 0), "null argument");
  DQNB_REQUIRE(n >= 0 && n < 65536 * 16, "n=%lld out of range", (long long)n);
  if (n == 0) return DQNB_OK;
  DeviceGuard guard(h->device);
  if (n <= kInlineSlots && h->S == 4 && h->plane_bytes % 16 == 0) {   // slot list by value, no upload
    DQNB_REQUIRE(s_tm1_dev && s_t_dev && a_dev && r_dev && disc_dev, "null argument");
    InlineSlots inl{};
    for (int64_t i = 0; i < n; ++i) {
      DQNB_REQUIRE(slots_host[i] >= 0 && slots_host[i] < h->capacity, "slot %lld outside [0,%lld)",
                   (long long)slots_host[i], (long long)h->capacity);
      inl.v[i] = (int32_t)slots_host[i];
    }
    cudaStream_t st = as_stream(stream);
    DQNB_TRY(flush(h, st));
    return launch_gather(h, nullptr, n, s_tm1_dev, s_t_dev, a_dev, r_dev, disc_dev, st, &inl);
  }
  const int32_t* dslots = nullptr;
  DQNB_TRY(upload_slots(h, slots_host, n, as_stream(stream), &dslots));
  return dqnb_replay_gather_dev(h, dslots, n, s_tm1_dev, s_t_dev, a_dev, r_dev, disc_dev, stream);
}

extern "C" int dqnb_replay_gather_host(dqnb_replay* h, const int64_t* slots_host, int64_t n,
                                       uint8_t* s_tm1_host, uint8_t* s_t_host, int32_t* a_host,
                                       float* r_host, float* disc_host, uintptr_t stream) {
  DQNB_REQUIRE(h && (n == 0 || (slots_host && s_tm1_host && s_t_host && a_host && r_host && disc_host)),
               "null argument");
  if (n == 0) return DQNB_OK;
  DeviceGuard guard(h->device);
  cudaStream_t st = as_stream(stream);
  const size_t stack_bytes = (size_t)h->plane_bytes * h->S;
  const size_t row = 2 * stack_bytes + 16;
  if ((int64_t)(row * n) > h->out_cap) {
    DQNB_CUDA(cudaDeviceSynchronize());
    cudaFree(h->d_out);
    h->out_cap = (int64_t)(row * n);
    DQNB_CUDA(cudaMalloc(&h->d_out, (size_t)h->out_cap));
  }
  uint8_t* d1 = h->d_out;
  uint8_t* d2 = d1 + stack_bytes * n;
  int32_t* da = reinterpret_cast<int32_t*>(d2 + stack_bytes * n);
  float* dr = reinterpret_cast<float*>(da + n);
  float* dd = dr + n;
  DQNB_TRY(dqnb_replay_gather(h, slots_host, n, d1, d2, da, dr, dd, stream));
  DQNB_CUDA(cudaMemcpyAsync(s_tm1_host, d1, stack_bytes * n, cudaMemcpyDeviceToHost, st));
  DQNB_CUDA(cudaMemcpyAsync(s_t_host, d2, stack_bytes * n, cudaMemcpyDeviceToHost, st));
  DQNB_CUDA(cudaMemcpyAsync(a_host, da, sizeof(int32_t) * n, cudaMemcpyDeviceToHost, st));
  DQNB_CUDA(cudaMemcpyAsync(r_host, dr, sizeof(float) * n, cudaMemcpyDeviceToHost, st));
  DQNB_CUDA(cudaMemcpyAsync(disc_host, dd, sizeof(float) * n, cudaMemcpyDeviceToHost, st));
  DQNB_CUDA(cudaStreamSynchronize(st));
  return DQNB_OK;
}

extern "C" int dqnb_replay_set_gather_splits(dqnb_replay* h, int splits) {
  DQNB_REQUIRE(h && splits >= 0, "bad argument");
  h->gather_splits = splits;
  return DQNB_OK;
}

extern "C" int dqnb_replay_stats(dqnb_replay* h, int64_t out[8]) {
  DQNB_REQUIRE(h && out, "null argument");
  int64_t used = h->pool_planes - (int64_t)h->free_list.size() - (int64_t)h->pending_free.size();
  out[0] = used; out[1] = h->pool_planes; out[2] = h->n_puts; out[3] = h->n_uploaded;
  out[4] = h->n_chained; out[5] = h->n_growths;
  out[6] = h->pool_planes * h->plane_stride + h->capacity * (8 * 4 + 12);
  out[7] = 0;
  return DQNB_OK;
}

// ------------------------------------------------------------------------- checkpoint (N2)
// One self-describing blob: header (int64 fields) + host bookkeeping + slot tables + the plane pool, the device
// parts copied at PCIe rate in one pass each - what get_state / set_state of the replay classes carry instead
// of one gathered 56 KB item per slot (replay.py:222-235, replay_circular.py:132-146,643-660).
namespace {
constexpr int64_t kBlobMagic = 0x44514E4252504C59LL;   // "DQNBRPLY"
enum BlobField { F_MAGIC, F_VERSION, F_CAPACITY, F_H, F_W, F_S, F_STRIDE, F_POOL, F_NFREE, F_HAVE_LAST,
                 F_OFF_SLOT_IDS, F_OFF_ACTIONS, F_OFF_REWARDS, F_OFF_DISCOUNTS, F_OFF_REFCOUNT, F_OFF_FREE,
                 F_OFF_LAST_IDS, F_OFF_LAST_ST, F_OFF_PLANES, F_TOTAL, F_COUNTERS /* 4 */, F_FIELDS = F_COUNTERS + 4 };
inline int64_t align64(int64_t x) { return (x + 63) / 64 * 64; }

void blob_layout(const dqnb_replay* h, int64_t n_free, int64_t f[F_FIELDS]) {
  memset(f, 0, sizeof(int64_t) * F_FIELDS);
  f[F_MAGIC] = kBlobMagic; f[F_VERSION] = 1; f[F_CAPACITY] = h->capacity; f[F_H] = h->H; f[F_W] = h->W; f[F_S] = h->S;
  f[F_STRIDE] = h->plane_stride; f[F_POOL] = h->pool_planes; f[F_NFREE] = n_free; f[F_HAVE_LAST] = h->have_last ? 1 : 0;
  int64_t o = align64((int64_t)sizeof(int64_t) * F_FIELDS);
  f[F_OFF_SLOT_IDS] = o; o = align64(o + 4 * 8 * h->capacity);
  f[F_OFF_ACTIONS] = o; o = align64(o + 4 * h->capacity);
  f[F_OFF_REWARDS] = o; o = align64(o + 4 * h->capacity);
  f[F_OFF_DISCOUNTS] = o; o = align64(o + 4 * h->capacity);
  f[F_OFF_REFCOUNT] = o; o = align64(o + 4 * h->pool_planes);
  f[F_OFF_FREE] = o; o = align64(o + 4 * n_free);
  f[F_OFF_LAST_IDS] = o; o = align64(o + 4 * 8);
  f[F_OFF_LAST_ST] = o; o = align64(o + (int64_t)h->plane_bytes * h->S);
  f[F_OFF_PLANES] = o; o += h->pool_planes * h->plane_stride;
  f[F_TOTAL] = o;
  f[F_COUNTERS] = h->n_puts; f[F_COUNTERS + 1] = h->n_uploaded; f[F_COUNTERS + 2] = h->n_chained; f[F_COUNTERS + 3] = h->n_growths;
}
}  // namespace

extern "C" int dqnb_replay_export_size(dqnb_replay* h, int64_t* bytes) {
  DQNB_REQUIRE(h && bytes, "null argument");
  int64_t f[F_FIELDS];
  blob_layout(h, (int64_t)(h->free_list.size() + h->pending_free.size()), f);
  *bytes = f[F_TOTAL];
  return DQNB_OK;
}

extern "C" int dqnb_replay_export(dqnb_replay* h, void* host, int64_t bytes, uintptr_t stream) {
  DQNB_REQUIRE(h && host, "null argument");
  DeviceGuard guard(h->device);
  cudaStream_t st = as_stream(stream);
  DQNB_TRY(flush(h, st));                       // staged puts reach the device first
  DQNB_CUDA(cudaDeviceSynchronize());            // commits / gathers may be on the side or a caller's gather stream
  // every released plane is reusable once the device is idle
  h->free_list.insert(h->free_list.end(), h->pending_free.begin(), h->pending_free.end());
  h->pending_free.clear();
  int64_t f[F_FIELDS];
  blob_layout(h, (int64_t)h->free_list.size(), f);
  DQNB_REQUIRE(bytes == f[F_TOTAL], "export buffer is %lld bytes, the snapshot needs %lld", (long long)bytes,
               (long long)f[F_TOTAL]);
  uint8_t* out = static_cast<uint8_t*>(host);
  memcpy(out, f, sizeof(f));
  memcpy(out + f[F_OFF_SLOT_IDS], h->slot_ids.data(), 4 * 8 * (size_t)h->capacity);
  DQNB_CUDA(cudaMemcpy(out + f[F_OFF_ACTIONS], h->d_actions, 4 * (size_t)h->capacity, cudaMemcpyDeviceToHost));
  DQNB_CUDA(cudaMemcpy(out + f[F_OFF_REWARDS], h->d_rewards, 4 * (size_t)h->capacity, cudaMemcpyDeviceToHost));
  DQNB_CUDA(cudaMemcpy(out + f[F_OFF_DISCOUNTS], h->d_discounts, 4 * (size_t)h->capacity, cudaMemcpyDeviceToHost));
  memcpy(out + f[F_OFF_REFCOUNT], h->refcount.data(), 4 * (size_t)h->pool_planes);
  memcpy(out + f[F_OFF_FREE], h->free_list.data(), 4 * h->free_list.size());
  memcpy(out + f[F_OFF_LAST_IDS], h->last_ids, sizeof(h->last_ids));
  memcpy(out + f[F_OFF_LAST_ST], h->last_st.data(), h->last_st.size());
  DQNB_CUDA(cudaMemcpy(out + f[F_OFF_PLANES], h->d_planes, (size_t)(h->pool_planes * h->plane_stride), cudaMemcpyDeviceToHost));
  return DQNB_OK;
}

extern "C" int dqnb_replay_import(dqnb_replay* h, const void* host, int64_t bytes, uintptr_t stream) {
  DQNB_REQUIRE(h && host && bytes >= (int64_t)sizeof(int64_t) * F_FIELDS, "null or truncated snapshot");
  const uint8_t* in = static_cast<const uint8_t*>(host);
  int64_t f[F_FIELDS];
  memcpy(f, in, sizeof(f));
  DQNB_REQUIRE(f[F_MAGIC] == kBlobMagic && f[F_VERSION] == 1, "not a replay snapshot (magic/version)");
  DQNB_REQUIRE(f[F_TOTAL] == bytes, "snapshot is %lld bytes, header says %lld", (long long)bytes, (long long)f[F_TOTAL]);
  DQNB_REQUIRE(f[F_CAPACITY] == h->capacity && f[F_H] == h->H && f[F_W] == h->W && f[F_S] == h->S &&
                   f[F_STRIDE] == h->plane_stride,
               "snapshot geometry (capacity %lld, %lldx%lldx%lld) does not match this store (capacity %lld, %dx%dx%d)",
               (long long)f[F_CAPACITY], (long long)f[F_H], (long long)f[F_W], (long long)f[F_S], (long long)h->capacity,
               h->H, h->W, h->S);
  DeviceGuard guard(h->device);
  cudaStream_t st = as_stream(stream);
  DQNB_CUDA(cudaDeviceSynchronize());
  // drop whatever was staged or in flight: the snapshot replaces everything
  for (Stage& s : h->stage) { s.n_planes = s.n_records = 0; s.in_flight = false; }
  std::fill(h->slot_staged.begin(), h->slot_staged.end(), 0);
  if (f[F_POOL] != h->pool_planes) {
    uint8_t* np = nullptr;
    cudaError_t e = cudaMalloc(&np, (size_t)(f[F_POOL] * h->plane_stride));
    if (e != cudaSuccess) { set_error("plane pool of %lld planes: %s", (long long)f[F_POOL], cudaGetErrorString(e)); return DQNB_ENOMEM; }
    cudaFree(h->d_planes);
    h->d_planes = np;
    h->pool_planes = f[F_POOL];
  }
  h->refcount.assign(reinterpret_cast<const int32_t*>(in + f[F_OFF_REFCOUNT]),
                     reinterpret_cast<const int32_t*>(in + f[F_OFF_REFCOUNT]) + h->pool_planes);
  h->free_list.assign(reinterpret_cast<const int32_t*>(in + f[F_OFF_FREE]),
                      reinterpret_cast<const int32_t*>(in + f[F_OFF_FREE]) + f[F_NFREE]);
  h->pending_free.clear();
  memcpy(h->slot_ids.data(), in + f[F_OFF_SLOT_IDS], 4 * 8 * (size_t)h->capacity);
  memcpy(h->last_ids, in + f[F_OFF_LAST_IDS], sizeof(h->last_ids));
  memcpy(h->last_st.data(), in + f[F_OFF_LAST_ST], h->last_st.size());
  h->have_last = f[F_HAVE_LAST] != 0;
  h->n_puts = f[F_COUNTERS]; h->n_uploaded = f[F_COUNTERS + 1]; h->n_chained = f[F_COUNTERS + 2]; h->n_growths = f[F_COUNTERS + 3];
  DQNB_CUDA(cudaMemcpy(h->d_slot_ids, in + f[F_OFF_SLOT_IDS], 4 * 8 * (size_t)h->capacity, cudaMemcpyHostToDevice));
  DQNB_CUDA(cudaMemcpy(h->d_actions, in + f[F_OFF_ACTIONS], 4 * (size_t)h->capacity, cudaMemcpyHostToDevice));
  DQNB_CUDA(cudaMemcpy(h->d_rewards, in + f[F_OFF_REWARDS], 4 * (size_t)h->capacity, cudaMemcpyHostToDevice));
  DQNB_CUDA(cudaMemcpy(h->d_discounts, in + f[F_OFF_DISCOUNTS], 4 * (size_t)h->capacity, cudaMemcpyHostToDevice));
  DQNB_CUDA(cudaMemcpy(h->d_planes, in + f[F_OFF_PLANES], (size_t)(h->pool_planes * h->plane_stride), cudaMemcpyHostToDevice));
  h->have_gather = h->have_commit = false;
  (void)st;
  return DQNB_OK;
}

// ---- internal hooks (learner_internal.cuh)
namespace dqnb {
int replay_gather_into(dqnb_replay* h, const int64_t* slots_host, int64_t n, uint8_t* s_tm1, uint8_t* s_t,
                       int32_t* a, float* r, float* d, cudaStream_t st) {
  return dqnb_replay_gather(h, slots_host, n, s_tm1, s_t, a, r, d, reinterpret_cast<uintptr_t>(st));
}
int replay_frame_shape(const dqnb_replay* h, int* height, int* width, int* stack) {
  *height = h->H; *width = h->W; *stack = h->S;
  return DQNB_OK;
}
}  // namespace dqnb
