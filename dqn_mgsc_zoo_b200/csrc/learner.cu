// DQN / double-DQN learner step on the device: Nature-CNN forward for 2-3 nets, TD/loss,
// backward, centered RMSProp, target sync; captured once as a CUDA graph with parallel
// branches.  Replaces the jitted `update` of dqn/agent.py:109-119 / double_q/agent.py:85-121
// and the target sync of dqn/agent.py:155-156.  See include/dqnb200.h.
#include <cuda.h>   // CUtensorMap and its enums only: the encoder is looked up through the runtime (no libcuda link)
#include <stdlib.h>

#include <algorithm>
#include <chrono>
#include <vector>

#include "head_kernels.cuh"
#include "learner_internal.cuh"
#include "nature_tf32.cuh"
#include "gemm_tf32_stream.cuh"
#include "act_kernel.cuh"
#include "nccl_dyn.cuh"
#include "hvp_kernels.cuh"
#include "conv23_fused.cuh"

using namespace dqnb;

namespace {

// The layer functors with their shape constants visible to generic lambdas (enqueue_hvp).
template <class L, class TIn, int BM>
struct HvpConvFwd : ConvFwd<L, TIn, BM> {
  static constexpr int LAYER_OUT = L::HOUT * L::WOUT, LAYER_COUT = L::COUT, LAYER_KCOL = L::KCOL;
};
template <class L, class TIn>
struct HvpConvWgrad : ConvWgrad<L, TIn> {
  static constexpr int LAYER_OUT = L::HOUT * L::WOUT, LAYER_COUT = L::COUT, LAYER_KCOL = L::KCOL;
};
template <class L>
struct HvpConvDgrad : ConvDgrad<L> {
  static constexpr int LAYER_CIN = L::CIN, LAYER_COUT = L::COUT;
};

struct ArenaLayout {
  int64_t w1, b1, w2, b2, w3, b3, fw1, fb1, fw2, fb2, total;
};

ArenaLayout make_layout(int A, int shared_bias) {
  ArenaLayout L;
  int64_t o = 0;
  L.w1 = o; o += (int64_t)Conv1::KCOL * Conv1::COUT;
  L.b1 = o; o += Conv1::COUT;
  L.w2 = o; o += (int64_t)Conv2::KCOL * Conv2::COUT;
  L.b2 = o; o += Conv2::COUT;
  L.w3 = o; o += (int64_t)Conv3::KCOL * Conv3::COUT;
  L.b3 = o; o += Conv3::COUT;
  L.fw1 = o; o += (int64_t)kFc1In * kFc1Out;
  L.fb1 = o; o += kFc1Out;
  L.fw2 = o; o += (int64_t)kFc1Out * A;
  L.fb2 = o; o += shared_bias ? 1 : A;
  L.total = o;
  return L;
}

// Tensor map of a dense row-major fp32 / fp16 matrix [outer][inner] for tiled TMA loads of {box_inner, box_outer} boxes.
// cuTensorMapEncodeTiled is a pure host-side encoder; it is fetched with cudaGetDriverEntryPoint so that the
// library keeps its single dependency on the CUDA runtime.
template <class T>
int make_tmap_2d(const T* base, uint64_t inner, uint64_t outer, uint32_t box_inner, uint32_t box_outer,
                 CUtensorMapSwizzle swizzle, CUtensorMap* out) {
  static_assert(sizeof(T) == 4 || sizeof(T) == 2, "fp32 or fp16 operands");
  using EncodeFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  static EncodeFn encode = [] {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q = cudaDriverEntryPointSymbolNotFound;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess) {
      cudaGetLastError();
      fn = nullptr;
    }
    return reinterpret_cast<EncodeFn>(fn);
  }();
  if (encode == nullptr) { set_error("cuTensorMapEncodeTiled is not available from this driver"); return DQNB_ECUDA; }
  const cuuint64_t dims[2] = {inner, outer};
  const cuuint64_t strides[1] = {inner * sizeof(T)};
  const cuuint32_t box[2] = {box_inner, box_outer};
  const cuuint32_t elem[2] = {1, 1};
  const CUresult r = encode(out, sizeof(T) == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2,
                            const_cast<T*>(base), dims, strides, box, elem,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed (CUresult %d)", (int)r); return DQNB_ECUDA; }
  return DQNB_OK;
}

int pick_splits(size_t ctas, int K, int want_ctas = 2 * kNumSMs) {
  int kchunks = (K + kBK - 1) / kBK;
  int s = (int)((want_ctas + ctas - 1) / ctas);
  s = std::min(s, std::max(1, kchunks / 4));
  return std::max(1, s);
}

}  // namespace

struct dqnb_learner {
  dqnb_learner_config cfg;
  int device = 0;
  int B = 0, A = 0, nets = 2;
  ArenaLayout L;
  // parameter / optimiser arenas (flat fp32, oracle leaf order)
  float *online = nullptr, *target = nullptr, *mu = nullptr, *nu = nullptr, *grad = nullptr;
  float *online_r = nullptr, *target_r = nullptr;   // TF32 mode: weights rounded to tf32 (GEMM operands)
  // TF32 slab mode: tensor maps of the fc1 weight shadows (device copies; gemm_tf32.cuh B_TMA).  [0] / [1]: online /
  // target as the forward's MN-major B operand, [2]: online as the dgrad's K-major B operand.  Null: cp.async staging.
  // [3]: both nets' fp16 conv1 filters (64 rows of 256 halves), [4] / [5]: online / target fp16 conv2 filters,
  // [6] / [7]: online / target fp16 conv3 filters - K-major B operands of the fp16 forward GEMMs;
  // [8] / [9]: online / target fp16 fc1 weights [3136][512] as the MN-major B operand of the fp16 fc1 forward.
  CUtensorMap* fc1_maps = nullptr;
  const void* fc1_fwd_map(const float* wts) const {
    return fc1_maps == nullptr ? nullptr : wts == online_r ? fc1_maps + 0 : wts == target_r ? fc1_maps + 1 : nullptr;
  }
  const void* h16_map(int slot) const { return fc1_maps == nullptr || !h16_maps ? nullptr : fc1_maps + slot; }
  bool h16_maps = false;
  __half* conv1_h16 = nullptr;   // TF32 mode, small batches: fp16 K-major conv1 filters [online 32 | target 32][256]
  // TF32 mode, small batches: fp16 K-major conv2 / conv3 filters, [online: conv2 64x512 | conv3 64x576][target: same],
  // and fp16 copies of the conv1 / conv2 outputs of every net (A operands of the fp16 conv2 / conv3 forward)
  __half* conv23_h16 = nullptr;
  __half *a1h[kMaxNets] = {}, *a2h[kMaxNets] = {};
  // fp16 fc1 forward (slab mode): plain fp16 copies of fc1's weights [online | target][3136][512] and of conv3's output
  __half* fc1_h16 = nullptr;
  __half* a3h[kMaxNets] = {};
  static constexpr size_t kFc1H = (size_t)kFc1In * kFc1Out;
  __half* fc1h(int target) const { return fc1_h16 + (target ? kFc1H : 0); }
  static constexpr size_t kC2H = (size_t)Conv2::COUT * Conv2::KCOL, kC3H = (size_t)Conv3::COUT * Conv3::KCOL;
  __half* c2h(int target) const { return conv23_h16 + (target ? kC2H + kC3H : 0); }
  __half* c3h(int target) const { return c2h(target) + kC2H; }
  H16Shadow h16_online() const {
    H16Shadow s = h16_none();
    if (conv1_h16) { s.dst[0] = conv1_h16; s.off[0] = L.w1; s.K[0] = Conv1::KCOL; s.N[0] = Conv1::COUT; }
    if (conv23_h16) {
      s.dst[1] = c2h(0); s.off[1] = L.w2; s.K[1] = Conv2::KCOL; s.N[1] = Conv2::COUT;
      s.dst[2] = c3h(0); s.off[2] = L.w3; s.K[2] = Conv3::KCOL; s.N[2] = Conv3::COUT;
    }
    if (fc1_h16) { s.plain_dst = fc1h(0); s.plain_off = L.fw1; s.plain_count = (int64_t)kFc1H; }
    return s;
  }
  // batch
  uint8_t *s_tm1 = nullptr, *s_t = nullptr;
  int32_t* act = nullptr;
  float *rew = nullptr, *disc = nullptr, *row_weight = nullptr;
  // activations per net, gradients
  float *a1[kMaxNets] = {}, *a2[kMaxNets] = {}, *a3[kMaxNets] = {}, *h1[kMaxNets] = {}, *q[kMaxNets] = {};
  float *dh1 = nullptr, *dy3 = nullptr, *dy2 = nullptr, *dy1 = nullptr;
  float *td = nullptr, *row_loss = nullptr, *dq_taken = nullptr, *loss = nullptr;
  // split-K scratch (one region per op so graph branches can overlap)
  float* ws = nullptr;
  unsigned* counters = nullptr;
  size_t ws_off = 0, cnt_off = 0, ws_cap = 0, cnt_cap = 0;
  bool sizing = false;
  // TF32 mode: split-K partials of the weight gradients, reduced inside the RMSProp kernel
  bool tf32 = false;
  // profiling (dqnb_learner_profile): one event after every launch, all work on one stream
  bool profiling = false;
  std::vector<cudaEvent_t> prof_ev;
  std::vector<const char*> prof_name;
  size_t prof_n = 0;
  long long* dbg_clocks = nullptr;   // [64][8] phase clocks of CTA 0 of each tensor-core launch
  // timeline (dqnb_learner_timeline): every launch of the update owns a slot {min begin, max end} in ns
  unsigned long long* d_tl = nullptr;
  std::vector<const char*> tl_name;
  int tl_n = 0, tl_graph_n = 0;
  bool tl_on = false;   // only launches of an update own timeline slots (q_values reuses the kernels)
  unsigned long long* next_tl(const char* name) {
    if (sizing || !tl_on || tl_n >= 64) return nullptr;
    if ((int)tl_name.size() <= tl_n) tl_name.push_back(name);
    tl_name[tl_n] = name;
    return d_tl ? d_tl + kTlSlot * tl_n++ : nullptr;
  }
  float* gpart = nullptr;
  size_t part_off = 0, part_cap = 0;
  float* fc1_part = nullptr;     // fc1 forward split-K partials [nets][splits][rows][512], summed by the head
  size_t fc1_part_cap = 0;
  int fc1_splits = 0, fc1_rows = 0;
  GradRegions regions;
  // Second batch-buffer set + its own graph (small batches): update_from_replay alternates between
  // the two, so the gather of update n+1 runs on `gather_st` beside update n instead of after it.
  uint8_t *alt_s_tm1 = nullptr, *alt_s_t = nullptr;
  int32_t* alt_act = nullptr;
  float *alt_rew = nullptr, *alt_disc = nullptr;
  cudaGraphExec_t graph_alt = nullptr;
  cudaStream_t gather_st = nullptr;
  cudaEvent_t set_free[2] = {}, gathered = nullptr;   // set_free[k]: last reader of buffer set k has finished
  bool set_used[2] = {false, false};
  int flip = 0;
  void swap_sets() {
    std::swap(s_tm1, alt_s_tm1); std::swap(s_t, alt_s_t); std::swap(act, alt_act);
    std::swap(rew, alt_rew); std::swap(disc, alt_disc);
  }
  // Acting (q_values of ONE observation with the online net, every environment step): its own small
  // graph - pinned observation -> s_tm1, the batch-1 forward, q -> pinned - built on first use.
  cudaGraphExec_t act_graph = nullptr;
  uint8_t* act_obs = nullptr;    // pinned [84*84*4]
  float* act_q = nullptr;        // pinned + device-mapped [kMaxActions]: the kernel writes q straight into it
  float* act_part = nullptr;     // fc1 partial sums of the acting kernel
  unsigned* act_seq = nullptr;   // device: acting launches executed
  unsigned* act_done = nullptr;  // pinned + mapped: sequence number of the last finished acting launch
  unsigned act_launched = 0;
  GridBarrier* act_bar = nullptr;
  // graph
  cudaGraphExec_t graph = nullptr;
  cudaStream_t side[3] = {};
  cudaStream_t chain = nullptr;   // stream of the dependent kernel chain in the enqueue in progress
  int prio_hi = 0;                // greatest launch priority of the device
  int prio_boost = 0;             // priority of the long side-branch kernels (see `boost`)
  bool pdl = false;               // programmatic dependent launch between consecutive chain kernels
  bool boost = false;             // next side-branch launch gets the chain's priority
  bool stream = false;            // large batch: GEMMs go to the persistent ring kernel (gemm_tf32_stream.cuh)
  cudaEvent_t ev[8] = {};
  int64_t updates = 0;
  int64_t launches_per_update = 0;
  // Hessian-vector product (dqnb_learner_hvp): tangent activations / cotangents, two temporaries per direction,
  // two gradient-shaped arenas and its own split-K scratch (the update graphs keep theirs); allocated on first use.
  struct Hvp {
    float *x1d = nullptr, *x2d = nullptr, *x3d = nullptr, *h1d = nullptr;      // tangent activations
    float *dh1d = nullptr, *dy3d = nullptr, *dy2d = nullptr, *dy1d = nullptr;  // tangent cotangents
    float *tu = nullptr, *tv = nullptr, *dqd = nullptr, *zero = nullptr;       // temporaries, dq tangent, zero bias
    float *ga = nullptr, *gb = nullptr;                                        // gradient-shaped partial results
    float* ws = nullptr; unsigned* counters = nullptr;
    bool ready = false;
  } hvp;
  // Data-parallel mode (dqnb_dp_init): the update graph itself all-reduces the gradient arena over NCCL - the
  // fc1 + fc2 part as soon as their weight gradients exist, the convolution part at the end - and then applies
  // the optimiser step, so one graph launch per rank is the whole data-parallel update.
  NcclComm dp_comm = nullptr;
  int dp_rank = 0, dp_world = 1;
  cudaStream_t dp_stream = nullptr;      // carries the collectives (high priority: NCCL's few CTAs must not queue behind whole GEMM grids)
  cudaEvent_t dp_ev[4] = {};
  bool dp_on() const { return dp_comm != nullptr; }
  // pinned staging for host-batch updates and slot uploads
  uint8_t* pinned = nullptr;
  size_t pinned_bytes = 0;
  int32_t* d_slots = nullptr;
  cudaEvent_t pinned_free = nullptr;
  bool pinned_busy = false;
  float* pinned_loss = nullptr;
  float* pinned_w = nullptr;                 // [kWRing][B] row weights on their way to the device (weighted updates)
  cudaEvent_t w_ev[4] = {};
  bool w_busy[4] = {};
  int w_cur = 0;
  // loss read-back without copies: the update writes (seq, loss) into a pinned ring (LossRing)
  unsigned long long* loss_ring = nullptr;   // pinned, device-mapped, kLossRing entries
  unsigned* d_seq = nullptr;                 // device: updates executed so far
  uint32_t seq_host = 0;                     // host: updates launched so far
  uint32_t loss_want[kLossRing] = {};        // sequence numbers of the outstanding posts
  uint64_t loss_posted = 0, loss_taken = 0;

  SplitK take_scratch(size_t ws_floats, size_t tiles) {
    SplitK sk{ws ? ws + ws_off : nullptr, counters ? counters + cnt_off : nullptr};
    ws_off += ws_floats;
    cnt_off += tiles;
    return sk;
  }
  // partial-gradient buffer for one wgrad op; registers the arena region it reduces into
  // `splits` partials of `stride` floats each; the first `count` of every partial belong to the arena
  // range [arena_off, arena_off + count), which the optimiser kernel sums in split order.
  float* take_part(int64_t arena_off, int64_t count, int splits, int64_t stride = 0) {
    if (stride == 0) stride = count;
    float* ptr = gpart ? gpart + part_off : nullptr;
    part_off += (size_t)stride * splits;
    GradRegion& r = regions.r[regions.n++];
    r.off = arena_off; r.count = count; r.part = ptr; r.stride = stride; r.splits = splits;
    return ptr;
  }
};

namespace {

// profiling hook: timestamp after the launch that was just enqueued on `st`
int mark(dqnb_learner* h, cudaStream_t st, const char* name) {
  if (!h->profiling) return DQNB_OK;
  if (h->prof_n == h->prof_ev.size()) {
    cudaEvent_t e;
    DQNB_CUDA(cudaEventCreate(&e));
    h->prof_ev.push_back(e);
    h->prof_name.push_back(name);
  }
  h->prof_name[h->prof_n] = name;
  DQNB_CUDA(cudaEventRecord(h->prof_ev[h->prof_n++], st));
  return DQNB_OK;
}

template <class P>
int run_gemm(dqnb_learner* h, const P& p, int Z, int splits, cudaStream_t st) {
  SplitK sk = h->take_scratch(gemm_ws_floats(p, Z, splits), gemm_tiles(p, Z));
  h->launches_per_update++;
  if (h->sizing) return DQNB_OK;
  sk.tl = h->next_tl(P::NAME);
  DQNB_CUDA(launch_gemm(p, Z, splits, sk, st));
  return mark(h, st, P::NAME);
}

template <class L, class TIn, int BM>
int conv_fwd(dqnb_learner* h, int nets, int rows, const TIn* const* x, const float* const* params,
             int64_t w_off, int64_t b_off, float* const* y, cudaStream_t st, bool linear = false) {
  ConvFwd<L, TIn, BM> p;
  p.linear = linear;
  p.M = rows * L::HOUT * L::WOUT; p.N = L::COUT; p.K = L::KCOL;
  for (int z = 0; z < nets; ++z) {
    p.x[z] = x[z]; p.w[z] = params[z] + w_off; p.bias[z] = params[z] + b_off; p.y[z] = y[z];
  }
  return run_gemm(h, p, nets, pick_splits(gemm_tiles(p, nets), p.K), st);
}

// Forward of `nets` networks over `rows` images each: x[z] uint8 NHWC -> h1[z], (q via head).
int forward_torso(dqnb_learner* h, int nets, int rows, const uint8_t* const* x, const float* const* params,
                  cudaStream_t st) {
  const float* in2[kMaxNets]; const float* in3[kMaxNets]; const float* in4[kMaxNets];
  for (int z = 0; z < nets; ++z) { in2[z] = h->a1[z]; in3[z] = h->a2[z]; in4[z] = h->a3[z]; }
  DQNB_TRY((conv_fwd<Conv1, uint8_t, 64>(h, nets, rows, x, params, h->L.w1, h->L.b1, h->a1, st)));
  DQNB_TRY((conv_fwd<Conv2, float, 32>(h, nets, rows, in2, params, h->L.w2, h->L.b2, h->a2, st)));
  DQNB_TRY((conv_fwd<Conv3, float, 32>(h, nets, rows, in3, params, h->L.w3, h->L.b3, h->a3, st)));
  DenseFwd p;
  p.M = rows; p.N = kFc1Out; p.K = kFc1In;
  for (int z = 0; z < nets; ++z) {
    p.x[z] = in4[z]; p.w[z] = params[z] + h->L.fw1; p.bias[z] = params[z] + h->L.fb1; p.y[z] = h->h1[z];
  }
  return run_gemm(h, p, nets, pick_splits(gemm_tiles(p, nets), p.K), st);
}

HeadArgs head_args(dqnb_learner* h, int rows, const float* const* params, int nets) {
  HeadArgs g;
  g.batch = rows; g.num_actions = h->A; g.shared_bias = h->cfg.shared_bias; g.double_q = nets == 3;
  g.grad_bound = h->cfg.grad_error_bound;
  g.inv_mean = 1.0f / (float)(h->cfg.loss_mean_over > 0 ? h->cfg.loss_mean_over : rows);
  for (int z = 0; z < kMaxNets; ++z) {
    int zz = std::min(z, nets - 1);
    g.h1[z] = h->h1[zz]; g.w2[z] = params[zz] + h->L.fw2; g.b2[z] = params[zz] + h->L.fb2; g.q[z] = h->q[zz];
  }
  g.fc1_part = nullptr; g.fc1_splits = 0; g.tl = nullptr; g.round_tf32 = 0;
  g.weight_inside = h->cfg.row_weighted == 2;
  for (int z = 0; z < kMaxNets; ++z) {
    int zz = std::min(z, nets - 1);
    g.b1[z] = params[zz] + h->L.fb1; g.h1_out[z] = h->h1[zz];
  }
  g.action = h->act; g.reward = h->rew; g.discount = h->disc; g.row_weight = h->row_weight;   // null unless cfg.row_weighted
  g.td = h->td; g.row_loss = h->row_loss; g.dq_taken = h->dq_taken; g.dh1 = h->dh1;
  return g;
}

}  // namespace

namespace {

template <class L, class TIn>
int conv_wgrad(dqnb_learner* h, int rows, const TIn* x, const float* dy, int64_t w_off, cudaStream_t st) {
  ConvWgrad<L, TIn> p;
  p.M = L::KCOL + 1; p.N = L::COUT; p.K = rows * L::HOUT * L::WOUT;
  p.x = x; p.dy = dy; p.grad = h->grad + w_off;
  return run_gemm(h, p, 1, pick_splits(gemm_tiles(p, 1), p.K), st);
}

template <class L>
int conv_dgrad(dqnb_learner* h, int rows, const float* dy, const float* w, const float* act_in, float* dx,
               cudaStream_t st) {
  ConvDgrad<L> p;
  p.M = rows * ConvDgrad<L>::NY * ConvDgrad<L>::NX; p.N = L::CIN;
  p.K = ConvDgrad<L>::T * ConvDgrad<L>::T * L::COUT;
  p.dy = dy; p.w = w; p.act_in = act_in; p.dx = dx;
  const int Z = L::STRIDE * L::STRIDE;
  return run_gemm(h, p, Z, pick_splits(gemm_tiles(p, Z), p.K), st);
}

// fork: `to` waits for everything recorded on `from` so far (works inside stream capture).
int fork(cudaEvent_t ev, cudaStream_t from, cudaStream_t to) {
  DQNB_CUDA(cudaEventRecord(ev, from));
  DQNB_CUDA(cudaStreamWaitEvent(to, ev, 0));
  return DQNB_OK;
}

// Enqueues one full update on `st` (+ side streams for independent branches).  In sizing mode
// nothing is launched; scratch offsets and the launch count are tallied instead.
int enqueue_update(dqnb_learner* h, cudaStream_t st) {
  h->ws_off = h->cnt_off = 0;
  h->tl_n = 0;
  h->launches_per_update = 0;
  const int B = h->B, nets = h->nets;
  const bool live = !h->sizing;
  cudaStream_t s1 = h->profiling ? st : h->side[0], s2 = h->profiling ? st : h->side[1], s3 = h->profiling ? st : h->side[2];
  const uint8_t* xin[kMaxNets] = {h->s_tm1, h->s_t, h->s_t};
  const float* params[kMaxNets] = {h->online, h->target, h->online};

  DQNB_TRY(forward_torso(h, nets, B, xin, params, st));
  HeadArgs g = head_args(h, B, params, nets);
  h->launches_per_update++;
  if (live) {
    g.tl = h->next_tl("head_td");
    head_td_for(B)<<<B, kHeadThreads, 0, st>>>(g);
    DQNB_CUDA(cudaGetLastError());
    DQNB_TRY(mark(h, st, "head_td"));
  }
  // branch s1: fc2 parameter grads + loss.   branch s2: fc1 weight grad.
  if (live) { DQNB_TRY(fork(h->ev[0], st, s1)); DQNB_TRY(fork(h->ev[0], st, s2)); }
  h->launches_per_update++;
  if (live) {
    int n = kFc1Out * h->A;
    fc2_wgrad_kernel<<<(n + 255) / 256, 256, 0, s1>>>(B, h->A, h->cfg.shared_bias, g.inv_mean, h->h1[0], h->act,
                                                     h->dq_taken, h->row_loss, h->grad + h->L.fw2,
                                                     h->grad + h->L.fb2, h->loss, LossRing{h->loss_ring, h->d_seq},
                                                     h->next_tl("fc2_wgrad"));
    DQNB_CUDA(cudaGetLastError());
    DQNB_TRY(mark(h, s1, "fc2_wgrad"));
  }
  {
    DenseWgrad p;
    p.M = kFc1In + 1; p.N = kFc1Out; p.K = B;
    p.x = h->a3[0]; p.dy = h->dh1; p.grad = h->grad + h->L.fw1;
    DQNB_TRY(run_gemm(h, p, 1, 1, s2));
  }
  {  // main: fc1 dgrad -> dy3 (masked by conv3's relu)
    DenseDgrad p;
    p.M = B; p.N = kFc1In; p.K = kFc1Out;
    p.dy = h->dh1; p.w = h->online + h->L.fw1; p.act = h->a3[0]; p.dx = h->dy3;
    DQNB_TRY(run_gemm(h, p, 1, pick_splits(gemm_tiles(p, 1), p.K), st));
  }
  if (live) DQNB_TRY(fork(h->ev[1], st, s3));
  DQNB_TRY((conv_wgrad<Conv3, float>(h, B, h->a2[0], h->dy3, h->L.w3, s3)));
  DQNB_TRY((conv_dgrad<Conv3>(h, B, h->dy3, h->online + h->L.w3, h->a2[0], h->dy2, st)));
  if (live) DQNB_TRY(fork(h->ev[2], st, s3));
  DQNB_TRY((conv_wgrad<Conv2, float>(h, B, h->a1[0], h->dy2, h->L.w2, s3)));
  DQNB_TRY((conv_dgrad<Conv2>(h, B, h->dy2, h->online + h->L.w2, h->a1[0], h->dy1, st)));
  DQNB_TRY((conv_wgrad<Conv1, uint8_t>(h, B, h->s_tm1, h->dy1, h->L.w1, st)));
  // join, then the optimiser over the whole arena
  if (live) {
    DQNB_TRY(fork(h->ev[3], s1, st));
    DQNB_TRY(fork(h->ev[4], s2, st));
    DQNB_TRY(fork(h->ev[5], s3, st));
  }
  if (h->cfg.defer_apply && !h->dp_on()) return DQNB_OK;
  if (h->dp_on()) {   // fp32 mode: one collective over the whole arena (the tf32 path chunks it)
    h->launches_per_update++;
    if (live) {
      const NcclApi* nccl = nccl_api();
      DQNB_REQUIRE(nccl != nullptr, "NCCL went away");
      DQNB_NCCL(nccl, nccl->AllReduce(h->grad, h->grad, (size_t)h->L.total, kNcclFloat32, kNcclSum, h->dp_comm, st));
    }
  }
  h->launches_per_update++;
  if (live) {
    int64_t n4 = (h->L.total >> 2) + 1;
    rmsprop_kernel<<<(unsigned)((n4 + 255) / 256), 256, 0, st>>>(h->online, h->mu, h->nu, h->grad, h->L.total,
                                                               h->cfg.learning_rate, h->cfg.decay, h->cfg.eps,
                                                               h->next_tl("rmsprop"));
    DQNB_CUDA(cudaGetLastError());
    DQNB_TRY(mark(h, st, "rmsprop"));
  }
  return DQNB_OK;
}

// ------------------------------------------------------------------ TF32 tensor-core path
constexpr int kStreamMinBatch = 256;   // below this the one-tile-per-CTA slab kernel (latency) wins
template <class P>
int run_t32(dqnb_learner* h, const P& p, int Z, int splits, cudaStream_t st) {
  SplitK sk = h->take_scratch(gemm_tf32_ws_floats(p, Z, splits), gemm_tf32_tiles(p, Z));
  h->launches_per_update++;
  if (h->sizing) return DQNB_OK;
  sk.tl = h->next_tl(P::NAME);
  if (h->profiling && getenv("DQNB_PROFILE_TWICE")) {   // warm the instruction cache with an untimed twin launch
    DQNB_CUDA(launch_gemm_tf32(p, Z, splits, sk, st));
    DQNB_TRY(mark(h, st, "(twin)"));
  }
  if (h->profiling && h->dbg_clocks) sk.dbg = h->dbg_clocks + 8 * h->prof_n;
  const bool on_chain = st == h->chain;
  if (h->stream) DQNB_CUDA(launch_gemm_tf32_stream(p, Z, splits, sk, st, on_chain ? h->prio_hi : 0));
  else
    DQNB_CUDA(launch_gemm_tf32(p, Z, splits, sk, st, on_chain ? h->prio_hi : h->boost ? h->prio_boost : 0,
                               on_chain && !h->profiling && h->pdl));
  return mark(h, st, P::NAME);
}

// Split-K factor of a TF32 GEMM: latency mode spreads one short tile's K over a cluster; stream mode
// splits only deferred-reduction problems, just enough to occupy every SM.
template <class P>
int t32_splits(const dqnb_learner* h, const P& p, int Z, int want_ctas = 120) {
  return h->stream ? t32s_splits_for(p, Z) : t32_splits_for(p, Z, want_ctas);
}

template <class L, class TIn>
int t32_conv_fwd(dqnb_learner* h, int nets, int rows, const TIn* const* x, const float* const* params,
                 const float* const* wts, int64_t w_off, int64_t b_off, float* const* y, cudaStream_t st,
                 __half* const* yh = nullptr) {
  T32ConvFwd<L, TIn> p;
  p.M = rows * L::HOUT * L::WOUT; p.N = L::COUT; p.K = L::KCOL;
  for (int z = 0; z < kMaxNets; ++z) {
    int zz = std::min(z, nets - 1);
    p.x[z] = x[zz]; p.w[z] = wts[zz] + w_off; p.bias[z] = params[zz] + b_off; p.y[z] = y[zz];
    p.yh[z] = yh ? yh[zz] : nullptr;
  }
  return run_t32(h, p, nets, t32_splits(h, p, nets), st);
}

// `params`: fp32 arenas (biases); `wts`: their tf32-rounded shadows (GEMM weight operands).
int forward_torso_tf32(dqnb_learner* h, int nets, int rows, const uint8_t* const* x, const float* const* params,
                       const float* const* wts, cudaStream_t st) {
  const float* in2[kMaxNets]; const float* in3[kMaxNets]; const float* in4[kMaxNets];
  for (int z = 0; z < nets; ++z) { in2[z] = h->a1[z]; in3[z] = h->a2[z]; in4[z] = h->a3[z]; }
  // conv2 / conv3 with fp16 operands (slab kernel only): every conv1 / conv2 variant below also leaves an fp16 copy
  bool h16 = h->conv23_h16 != nullptr && !h->stream;
  for (int z = 0; z < nets; ++z) h16 = h16 && (wts[z] == h->online_r || wts[z] == h->target_r);
  // With fp16 operands downstream, the fp32 conv1 / conv2 outputs of the nets that are only evaluated (z >= 1: no
  // backward pass reads them) are never consumed: their stores (two thirds of the layer's fp32 bytes) are skipped.
  float* y1[kMaxNets]; float* y2[kMaxNets];
  for (int z = 0; z < kMaxNets; ++z) { y1[z] = (h16 && z > 0) ? nullptr : h->a1[z]; y2[z] = (h16 && z > 0) ? nullptr : h->a2[z]; }
  if (nets == 3 && x[1] == x[2] && wts[0] == wts[2]) {   // double-Q update: one merged conv1 GEMM
    auto fill = [&](auto& p) {
      p.Mh = rows * Conv1::HOUT * Conv1::WOUT;
      p.M = 2 * p.Mh; p.N = 2 * Conv1::COUT; p.K = Conv1::KCOL;
      p.x[0] = x[0]; p.x[1] = x[1];
      p.w[0] = wts[0] + h->L.w1; p.w[1] = wts[1] + h->L.w1;
      p.bias[0] = params[0] + h->L.b1; p.bias[1] = params[1] + h->L.b1;
      for (int z = 0; z < kMaxNets; ++z) { p.y[z] = y1[z]; p.yh[z] = h16 ? h->a1h[z] : nullptr; }
    };
    if (h->conv1_h16 != nullptr && wts[0] == h->online_r && wts[1] == h->target_r) {   // fp16 operands: one wave at batch 32
      H16ConvFwdDoubleQ<Conv1, uint8_t> p;
      fill(p);
      p.K = Conv1::KCOL / 2;   // 32-bit words
      p.w16 = h->conv1_h16;
      p.bmap = h->h16_map(3);
      DQNB_TRY(run_t32(h, p, 1, t32_splits(h, p, 1), st));
    } else {
      T32ConvFwdDoubleQ<Conv1, uint8_t> p;
      fill(p);
      DQNB_TRY(run_t32(h, p, 1, t32_splits(h, p, 1), st));
    }
  } else {
    DQNB_TRY((t32_conv_fwd<Conv1, uint8_t>(h, nets, rows, x, params, wts, h->L.w1, h->L.b1, y1, st,
                                           h16 ? h->a1h : nullptr)));
  }
  if (h16) {
    auto run_h16 = [&](auto p, __half* const* xin, __half* const* yh, float* const* y, int64_t b_off, int layer) -> int {
      using P = decltype(p);
      p.M = rows * P::HOUT_WOUT; p.N = P::BN; p.K = P::KWORDS;
      for (int z = 0; z < kMaxNets; ++z) {
        const int zz = std::min(z, nets - 1);
        const int tgt = wts[zz] == h->target_r;
        p.x[z] = xin[zz]; p.w16[z] = layer == 2 ? h->c2h(tgt) : h->c3h(tgt);
        p.bmap[z] = h->h16_map((layer == 2 ? 4 : 6) + tgt);
        p.bias[z] = params[zz] + b_off; p.y[z] = y[zz]; p.yh[z] = yh ? yh[zz] : nullptr;
      }
      // split-K over a cluster (DQNB_H16_SPLITS2 / 3, default 1 = the whole K range in one slab)
      static const int s2 = getenv("DQNB_H16_SPLITS2") ? atoi(getenv("DQNB_H16_SPLITS2")) : 1;
      static const int s3 = getenv("DQNB_H16_SPLITS3") ? atoi(getenv("DQNB_H16_SPLITS3")) : 1;
      return run_t32(h, p, nets, layer == 2 ? s2 : s3, st);
    };
    static const bool no_fused = getenv("DQNB_NO_FUSED_C23") != nullptr;
    if (!no_fused && h->h16_map(4) != nullptr) {
      // conv2 + conv3 in one kernel, one CTA per (image, net) (conv23_fused.cuh)
      Conv23Args g = {};
      for (int z = 0; z < kMaxNets; ++z) {
        const int zz = std::min(z, nets - 1);
        const int tgt = wts[zz] == h->target_r;
        g.a1h[z] = h->a1h[zz]; g.map2[z] = h->h16_map(4 + tgt); g.map3[z] = h->h16_map(6 + tgt);
        g.bias2[z] = params[zz] + h->L.b2; g.bias3[z] = params[zz] + h->L.b3;
        g.a2[z] = y2[zz]; g.a3[z] = h->a3[zz]; g.a3h[z] = h->a3h[zz];
      }
      h->launches_per_update++;
      if (!h->sizing) {
        g.tl = h->next_tl("conv23_fwd");
        g.dbg = (h->profiling && h->dbg_clocks) ? h->dbg_clocks + 8 * h->prof_n : nullptr;
        const bool on_chain = st == h->chain;
        DQNB_CUDA(launch_conv23_fwd(g, rows, nets, st, on_chain ? h->prio_hi : 0, on_chain && !h->profiling && h->pdl));
        DQNB_TRY(mark(h, st, "conv23_fwd"));
      }
    } else {
      DQNB_TRY(run_h16(H16ConvFwd<Conv2>{}, h->a1h, h->a2h, y2, h->L.b2, 2));
      DQNB_TRY(run_h16(H16ConvFwd<Conv3>{}, h->a2h, h->fc1_h16 != nullptr ? h->a3h : nullptr, h->a3, h->L.b3, 3));
    }
  } else {
    DQNB_TRY((t32_conv_fwd<Conv2, float>(h, nets, rows, in2, params, wts, h->L.w2, h->L.b2, h->a2, st)));
    DQNB_TRY((t32_conv_fwd<Conv3, float>(h, nets, rows, in3, params, wts, h->L.w3, h->L.b3, h->a3, st)));
  }
  if (h16 && h->fc1_h16 != nullptr && h->h16_map(8) != nullptr) {   // fp16 operands, weights by TMA only
    H16DenseFwd p;
    p.merged = nets == 3;
    p.batch = rows;
    p.M = p.merged ? 2 * rows : rows; p.N = kFc1Out; p.K = kFc1In / 2;
    const int Z = p.merged ? 2 : nets;
    for (int z = 0; z < kMaxNets; ++z) {
      const int zz = std::min(z, nets - 1);
      p.x[z] = h->a3h[zz]; p.bmap[z] = h->h16_map(8 + (wts[zz] == h->target_r));
    }
    p.splits = t32_splits(h, p, Z, kNumSMs);
    h->fc1_splits = p.splits;
    h->fc1_rows = rows;
    const size_t need = (size_t)nets * p.splits * rows * kFc1Out;
    if (h->sizing) h->fc1_part_cap = std::max(h->fc1_part_cap, need);
    p.part = h->fc1_part;
    return run_t32(h, p, Z, p.splits, st);
  }
  T32DenseFwd p;
  p.merged = nets == 3;
  p.batch = rows;
  p.M = p.merged ? 2 * rows : rows; p.N = kFc1Out; p.K = kFc1In;
  const int Z = p.merged ? 2 : nets;
  for (int z = 0; z < kMaxNets; ++z) {
    int zz = std::min(z, nets - 1);
    p.x[z] = in4[zz]; p.w[z] = wts[zz] + h->L.fw1;
    p.bmap[z] = h->stream ? nullptr : h->fc1_fwd_map(wts[zz]);
  }
  p.splits = t32_splits(h, p, Z, kNumSMs);
  h->fc1_splits = p.splits;
  h->fc1_rows = rows;
  const size_t need = (size_t)nets * p.splits * rows * kFc1Out;
  if (h->sizing) h->fc1_part_cap = std::max(h->fc1_part_cap, need);
  p.part = h->fc1_part;
  return run_t32(h, p, Z, p.splits, st);
}

// `bias_st`: stream for the separate bias column-sum kernel (BIAS_ROW = false); it must already
// be ordered after the producer of dy.
template <class L, class TIn, int KB, bool BIAS_ROW>
int t32_conv_wgrad(dqnb_learner* h, int rows, const TIn* x, const float* dy, int64_t w_off, cudaStream_t st,
                   cudaStream_t bias_st = nullptr) {
  using P = T32ConvWgrad<L, TIn, KB, BIAS_ROW>;
  P p;
  p.M = P::ROWS; p.N = L::COUT; p.K = rows * L::HOUT * L::WOUT;
  p.x = x; p.dy = dy;
  const int splits = t32_splits(h, p, 1);
  // Stream mode has a handful of long splits, far too few blocks for the bias column sums: those get
  // their own partial buffer (one row of COUT per block) and their own region.
  const bool own_bias = !BIAS_ROW && h->stream;
  p.gpart = h->take_part(w_off, own_bias ? (int64_t)L::KCOL * L::COUT : (int64_t)P::PART, splits, (int64_t)P::PART);
  if (!BIAS_ROW) {
    h->launches_per_update++;
    const int blocks = own_bias ? std::max(1, std::min(4 * kNumSMs, p.K / 64)) : splits;
    float* bpart = own_bias ? h->take_part(w_off + (int64_t)L::KCOL * L::COUT, L::COUT, blocks) : p.gpart;
    if (!h->sizing) {
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(blocks); cfg.blockDim = dim3(256); cfg.stream = bias_st;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributePriority;
      attr[0].val.priority = h->boost ? h->prio_boost : 0;
      cfg.attrs = attr; cfg.numAttrs = 1;
      DQNB_CUDA(cudaLaunchKernelEx(&cfg, bias_colsum_kernel<L::COUT>, dy, p.K, bpart,
                                   own_bias ? (size_t)L::COUT : (size_t)P::PART,
                                   own_bias ? (size_t)0 : (size_t)L::KCOL * L::COUT, h->next_tl("bias_colsum")));
      DQNB_TRY(mark(h, bias_st, "bias_colsum"));
    }
  }
  return run_t32(h, p, 1, splits, st);
}

template <class L, int KB>
int t32_conv_dgrad(dqnb_learner* h, int rows, const float* dy, const float* w, const float* act_in, float* dx,
                   cudaStream_t st) {
  T32ConvDgrad<L, KB> p;
  p.M = rows * T32ConvDgrad<L, KB>::NY * T32ConvDgrad<L, KB>::NX; p.N = L::CIN;
  p.K = T32ConvDgrad<L, KB>::T * T32ConvDgrad<L, KB>::T * L::COUT;
  p.dy = dy; p.w = w; p.act_in = act_in; p.dx = dx;
  const int Z = L::STRIDE * L::STRIDE;
  return run_t32(h, p, Z, t32_splits(h, p, Z), st);
}

template <int KB, int BN>
int t32_dense_wgrad(dqnb_learner* h, int rows, bool fuse, cudaStream_t st) {
  T32DenseWgrad<KB, BN> p;
  p.M = kFc1In + 1; p.N = kFc1Out; p.K = rows;
  p.x = h->a3[0]; p.dy = h->dh1;
  const int splits = t32_splits(h, p, 1);
  p.fuse = FusedRmsProp{nullptr, nullptr, nullptr, nullptr, nullptr, 0.f, 0.f, 0.f};
  p.gpart = nullptr;
  if (fuse && splits == 1) {
    const int64_t o = h->L.fw1;
    p.fuse = FusedRmsProp{h->online + o, h->online_r + o, h->mu + o, h->nu + o, h->grad + o,
                          h->cfg.learning_rate, h->cfg.decay, h->cfg.eps};
    p.fuse.p_h = h->fc1_h16;   // (online copy; null when the fp16 fc1 forward is off)
  } else {
    p.gpart = h->take_part(h->L.fw1, (int64_t)p.M * p.N, splits);
  }
  return run_t32(h, p, 1, splits, st);
}

// One full update with every GEMM on tcgen05 (TF32).  Same dependency structure as the fp32
// graph; the weight-gradient partials are summed by the optimiser kernel.
int enqueue_update_tf32(dqnb_learner* h, cudaStream_t st) {
  h->ws_off = h->cnt_off = h->part_off = 0;
  h->regions.n = 0;
  h->tl_n = 0;
  h->launches_per_update = 0;
  h->chain = st;
  const int B = h->B, nets = h->nets;
  const bool live = !h->sizing;
  cudaStream_t s1 = h->profiling ? st : h->side[0], s2 = h->profiling ? st : h->side[1], s3 = h->profiling ? st : h->side[2];
  const uint8_t* xin[kMaxNets] = {h->s_tm1, h->s_t, h->s_t};
  const float* params[kMaxNets] = {h->online, h->target, h->online};

  const float* wts[kMaxNets] = {h->online_r, h->target_r, h->online_r};
  DQNB_TRY(forward_torso_tf32(h, nets, B, xin, params, wts, st));
  HeadArgs g = head_args(h, B, params, nets);
  g.round_tf32 = 1;
  g.fc1_part = h->fc1_part; g.fc1_splits = h->fc1_splits;
  h->launches_per_update++;
  if (live) {
    g.tl = h->next_tl("head_td");
    // a link of the dependent chain: chain priority, and (programmatic dependent launch) placed while fc1 still
    // stores its partials - the kernel prefetches its own operands ahead of its griddepcontrol.wait
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(B); cfg.blockDim = dim3(kHeadThreads); cfg.stream = st;
    cudaLaunchAttribute attr[2];
    unsigned na = 0;
    attr[na].id = cudaLaunchAttributePriority;
    attr[na++].val.priority = h->prio_hi;
    if (h->pdl && !h->profiling && getenv("DQNB_NO_HEAD_PDL") == nullptr) {
      attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      attr[na++].val.programmaticStreamSerializationAllowed = 1;
    }
    cfg.attrs = attr; cfg.numAttrs = na;
    DQNB_CUDA(cudaLaunchKernelEx(&cfg, head_td_for(B), g));
    DQNB_TRY(mark(h, st, "head_td"));
  }
  if (live) DQNB_TRY(fork(h->ev[0], st, s1));
  h->launches_per_update++;
  float* fc2_part = nullptr;
  constexpr int kFc2Slices = 32;
  if (h->stream) fc2_part = h->take_part(h->L.fw2, (int64_t)kFc1Out * h->A, kFc2Slices);
  if (live) {
    int n = kFc1Out * h->A;
    if (h->stream)
      fc2_wgrad_sliced_kernel<<<dim3((n + 255) / 256, kFc2Slices), 256, 0, s1>>>(
          B, h->A, h->cfg.shared_bias, g.inv_mean, h->h1[0], h->act, h->dq_taken, h->row_loss, fc2_part,
          h->grad + h->L.fb2, h->loss, LossRing{h->loss_ring, h->d_seq}, h->next_tl("fc2_wgrad"));
    else
      fc2_wgrad_kernel<<<(n + 255) / 256, 256, 0, s1>>>(B, h->A, h->cfg.shared_bias, g.inv_mean, h->h1[0], h->act,
                                                       h->dq_taken, h->row_loss, h->grad + h->L.fw2,
                                                       h->grad + h->L.fb2, h->loss, LossRing{h->loss_ring, h->d_seq},
                                                       h->next_tl("fc2_wgrad"));
    DQNB_CUDA(cudaGetLastError());
    DQNB_TRY(mark(h, s1, "fc2_wgrad"));
  }
  {
    T32DenseDgrad p;
    p.M = B; p.N = kFc1In; p.K = kFc1Out;
    p.dy = h->dh1; p.w = h->online_r + h->L.fw1; p.act = h->a3[0]; p.dx = h->dy3;
    p.bmap = h->fc1_maps != nullptr && !h->stream ? h->fc1_maps + 2 : nullptr;
    DQNB_TRY(run_t32(h, p, 1, t32_splits(h, p, 1), st));
  }
  static const bool fc1w_early = getenv("DQNB_FC1W_EARLY") != nullptr;   // release fc1's wgrad right behind its dgrad
  if (live) DQNB_TRY(fork(h->ev[1], st, s3));
  if (live && fc1w_early) DQNB_TRY(fork(h->ev[1], st, s2));
  DQNB_TRY((t32_conv_wgrad<Conv3, float, 8, true>(h, B, h->a2[0], h->dy3, h->L.w3, s3)));
  DQNB_TRY((t32_conv_dgrad<Conv3, 6>(h, B, h->dy3, h->online_r + h->L.w3, h->a2[0], h->dy2, st)));
  if (live) { DQNB_TRY(fork(h->ev[2], st, s3)); if (!fc1w_early) DQNB_TRY(fork(h->ev[2], st, s2)); DQNB_TRY(fork(h->ev[2], st, s1)); }
  // fc1's weight gradient is ordered after fc1's dgrad (which reads the old weights) so that, when
  // the batch is one K block (no split-K), its epilogue can apply RMSProp to fc1 on the spot.  It
  // is bandwidth-bound (7 streams over 1.6 M parameters) and is released here, where the chain
  // kernels (conv2 dgrad, conv1 wgrad) leave room for its small CTAs beside their own on every SM.
  // conv2's weight gradient: half slabs (88 CTAs of 97 KB) fit beside the CTAs of fc1's wgrad and conv1's wgrad, so
  // it overlaps both; as 44 CTAs of 193 KB (DQNB_C2W_KB8) each needed an SM to itself and the kernel sat at the very
  // end of the update (measured: span 72.3 -> 69.9 us).  DQNB_C2W_FIRST enqueues it ahead of fc1's wgrad (no gain).
  static const bool c2w_first = getenv("DQNB_C2W_FIRST") != nullptr, c2w_kb4 = getenv("DQNB_C2W_KB8") == nullptr;
  auto conv2_wgrad = [&]() -> int {
    h->boost = h->pdl;   // see the fc1 wgrad below: conv2's wgrad and both bias column sums would otherwise wait for the chain
    const int rc = c2w_kb4 ? t32_conv_wgrad<Conv2, float, 4, false>(h, B, h->a1[0], h->dy2, h->L.w2, s3, s1)
                           : t32_conv_wgrad<Conv2, float, 8, false>(h, B, h->a1[0], h->dy2, h->L.w2, s3, s1);
    h->boost = false;
    return rc;
  };
  if (c2w_first) DQNB_TRY(conv2_wgrad());
  bool fused = false;
  if (B <= kT32BK && !h->cfg.defer_apply && !h->stream) {
    fused = true;
    if (live && !h->profiling && !h->pdl && !getenv("DQNB_NO_STAGGER")) { stagger_kernel<<<1, 32, 0, s2>>>(); DQNB_CUDA(cudaGetLastError()); }
    // With dependent launches the chain always has CTAs pending, so a low-priority kernel would only
    // run after the chain; this one is the longest side branch (bandwidth-bound, ~10 us) and must
    // start as soon as its inputs exist.
    h->boost = h->pdl;
    int rc_w = t32_dense_wgrad<1, 32>(h, B, /*fuse=*/true, s2);
    h->boost = false;
    DQNB_TRY(rc_w);
  }
  else DQNB_TRY((t32_dense_wgrad<4, 128>(h, B, false, s2)));
  const int64_t dp_split_end = h->stream ? h->L.fb2 : h->L.fw2;
  const bool dp_live = live && h->dp_on();
  cudaStream_t sd = h->profiling ? st : h->dp_stream;
  if (h->dp_on()) h->launches_per_update += 2;
  if (dp_live) {
    // chunk 1 = [fc1.w, end of arena): both producers (fc1 wgrad on s2, fc2 wgrad first on s1) are enqueued;
    // sum their split-K partials and start the collective while conv2 dgrad / conv2 wgrad / conv1 wgrad run
    const NcclApi* nccl = nccl_api();
    DQNB_REQUIRE(nccl != nullptr, "NCCL went away");
    DQNB_TRY(fork(h->dp_ev[0], s1, sd));
    DQNB_TRY(fork(h->dp_ev[1], s2, sd));
    grad_reduce_range_kernel<<<grad_reduce_range_blocks(h->L.fw1, dp_split_end), 256, 0, sd>>>(
        h->grad, h->regions, h->L.fw1, dp_split_end, h->next_tl("grad_reduce_fc"));
    DQNB_CUDA(cudaGetLastError());
    DQNB_TRY(mark(h, sd, "grad_reduce_fc"));
    DQNB_NCCL(nccl, nccl->AllReduce(h->grad + h->L.fw1, h->grad + h->L.fw1, (size_t)(h->L.total - h->L.fw1),
                                    kNcclFloat32, kNcclSum, h->dp_comm, sd));
    DQNB_TRY(mark(h, sd, "allreduce_fc"));
  }
  if (!c2w_first) DQNB_TRY(conv2_wgrad());
  // slab mode: the whole K range (4 taps x 64 channels = 8 k-blocks) in one slab: 100 unsplit tiles, one wave, no
  // cluster reduce (as 2-CTA clusters it was 200 CTAs over two waves: 11.5 us); DQNB_C2DGRAD_KB4 restores that
  static const bool c2_kb4 = getenv("DQNB_C2DGRAD_KB4") != nullptr;
  if (h->stream || c2_kb4) DQNB_TRY((t32_conv_dgrad<Conv2, 4>(h, B, h->dy2, h->online_r + h->L.w2, h->a1[0], h->dy1, st)));
  else DQNB_TRY((t32_conv_dgrad<Conv2, 8>(h, B, h->dy2, h->online_r + h->L.w2, h->a1[0], h->dy1, st)));
  if (live) DQNB_TRY(fork(h->ev[6], st, s1));
  h->boost = h->pdl;
  int rc_c1 = t32_conv_wgrad<Conv1, uint8_t, 4, false>(h, B, h->s_tm1, h->dy1, h->L.w1, st, s1);
  h->boost = false;
  DQNB_TRY(rc_c1);
  if (live) {
    DQNB_TRY(fork(h->ev[3], s1, st));
    DQNB_TRY(fork(h->ev[4], s2, st));
    DQNB_TRY(fork(h->ev[5], s3, st));
  }
  if (h->dp_on()) {
    // chunk 2 = the convolution gradients [0, fc1.w): reduce, all-reduce behind chunk 1 on the collective
    // stream, then the optimiser step over the whole (now global) gradient
    h->launches_per_update += 3;
    if (live) {
      const NcclApi* nccl = nccl_api();
      DQNB_REQUIRE(nccl != nullptr, "NCCL went away");
      grad_reduce_range_kernel<<<grad_reduce_range_blocks(0, h->L.fw1), 256, 0, st>>>(
          h->grad, h->regions, 0, h->L.fw1, h->next_tl("grad_reduce_conv"));
      DQNB_CUDA(cudaGetLastError());
      DQNB_TRY(mark(h, st, "grad_reduce_conv"));
      DQNB_TRY(fork(h->dp_ev[2], st, sd));
      DQNB_NCCL(nccl, nccl->AllReduce(h->grad, h->grad, (size_t)h->L.fw1, kNcclFloat32, kNcclSum, h->dp_comm, sd));
      DQNB_TRY(mark(h, sd, "allreduce_conv"));
      DQNB_TRY(fork(h->dp_ev[3], sd, st));
      GradRegions none{};
      rmsprop_reduce_kernel<true><<<rmsprop_reduce_blocks(h->L.total, 0, 0, 0), 256, 0, st>>>(
          h->online, h->online_r, h->h16_online(), h->mu, h->nu, h->grad, h->L.total, none, 0, 0, 0,
          h->cfg.learning_rate, h->cfg.decay, h->cfg.eps, h->next_tl("rmsprop"));
      DQNB_CUDA(cudaGetLastError());
      DQNB_TRY(mark(h, st, "rmsprop"));
    }
    return DQNB_OK;
  }
  h->launches_per_update++;
  if (live) {
    const int64_t split_end = fused ? h->L.fw1 : (h->stream ? h->L.fb2 : h->L.fw2);   // stream: fc2's slices too
    const int64_t skip_lo = fused ? h->L.fw1 : 0, skip_hi = fused ? h->L.fw2 : 0;
    if (h->cfg.defer_apply)   // sum the split-K partials into DQNB_GRAD, nothing else
      rmsprop_reduce_kernel<false><<<rmsprop_reduce_blocks(h->L.total, split_end, skip_lo, skip_hi), 256, 0, st>>>(
          h->online, h->online_r, h16_none(), h->mu, h->nu, h->grad, h->L.total, h->regions, split_end, skip_lo, skip_hi,
          h->cfg.learning_rate, h->cfg.decay, h->cfg.eps, h->next_tl("grad_reduce"));
    else {
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(rmsprop_reduce_blocks(h->L.total, split_end, skip_lo, skip_hi)); cfg.blockDim = dim3(256); cfg.stream = st;
      cudaLaunchAttribute attr[2];
      unsigned na = 0;
      attr[na].id = cudaLaunchAttributePriority;
      attr[na++].val.priority = h->prio_hi;
      if (h->pdl && !h->profiling && getenv("DQNB_NO_RMSPROP_PDL") == nullptr) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na++].val.programmaticStreamSerializationAllowed = 1;
      }
      cfg.attrs = attr; cfg.numAttrs = na;
      DQNB_CUDA(cudaLaunchKernelEx(&cfg, rmsprop_reduce_kernel<true>, h->online, h->online_r, h->h16_online(), h->mu, h->nu,
                                   h->grad, h->L.total, h->regions, split_end, skip_lo, skip_hi, h->cfg.learning_rate,
                                   h->cfg.decay, h->cfg.eps, h->next_tl("rmsprop_reduce")));
    }
    DQNB_CUDA(cudaGetLastError());
    DQNB_TRY(mark(h, st, "rmsprop_reduce"));
  }
  return DQNB_OK;
}

int enqueue_any(dqnb_learner* h, cudaStream_t st) {
  h->tl_on = true;
  const int rc = h->tf32 ? enqueue_update_tf32(h, st) : enqueue_update(h, st);
  h->tl_on = false;
  return rc;
}

// MGSC adjoint dots: the four layer-local forward products of the direction `v` (an arena laid out
// like the parameters) on the activations the last update saved, into the dead target-net
// activation buffers, then one dot kernel.  fp32 SIMT GEMMs in both precision modes.
int enqueue_adjoint(dqnb_learner* h, const float* v, float* out, cudaStream_t st) {
  h->ws_off = h->cnt_off = 0;
  const int B = h->B;
  const uint8_t* x1[kMaxNets] = {h->s_tm1, nullptr, nullptr};
  const float* x2[kMaxNets] = {h->a1[0], nullptr, nullptr};
  const float* x3[kMaxNets] = {h->a2[0], nullptr, nullptr};
  const float* vp[kMaxNets] = {v, nullptr, nullptr};
  float* z1[kMaxNets] = {h->a1[1], nullptr, nullptr};
  float* z2[kMaxNets] = {h->a2[1], nullptr, nullptr};
  float* z3[kMaxNets] = {h->a3[1], nullptr, nullptr};
  // the four products are independent: one per stream (each op owns its slice of the split-K scratch)
  const bool live = !h->sizing;
  cudaStream_t s1 = live ? h->side[0] : st, s2 = live ? h->side[1] : st, s3 = live ? h->side[2] : st;
  if (live) {
    DQNB_CUDA(cudaEventRecord(h->ev[0], st));
    DQNB_CUDA(cudaStreamWaitEvent(s1, h->ev[0], 0));
    DQNB_CUDA(cudaStreamWaitEvent(s2, h->ev[0], 0));
    DQNB_CUDA(cudaStreamWaitEvent(s3, h->ev[0], 0));
  }
  DQNB_TRY((conv_fwd<Conv1, uint8_t, 64>(h, 1, B, x1, vp, h->L.w1, h->L.b1, z1, st, true)));
  DQNB_TRY((conv_fwd<Conv2, float, 32>(h, 1, B, x2, vp, h->L.w2, h->L.b2, z2, s1, true)));
  DQNB_TRY((conv_fwd<Conv3, float, 32>(h, 1, B, x3, vp, h->L.w3, h->L.b3, z3, s2, true)));
  DenseFwd p;
  p.linear = 1;
  p.M = B; p.N = kFc1Out; p.K = kFc1In;
  p.x[0] = h->a3[0]; p.w[0] = v + h->L.fw1; p.bias[0] = v + h->L.fb1; p.y[0] = h->h1[1];
  DQNB_TRY(run_gemm(h, p, 1, pick_splits(gemm_tiles(p, 1), p.K), s3));
  if (h->sizing) return DQNB_OK;
  DQNB_TRY(fork(h->ev[3], s1, st));
  DQNB_TRY(fork(h->ev[4], s2, st));
  DQNB_TRY(fork(h->ev[5], s3, st));
  AdjointArgs g;
  g.num_actions = h->A; g.shared_bias = h->cfg.shared_bias;
  g.dy1 = h->dy1; g.z1 = h->a1[1]; g.n1 = Conv1::OUT_PER_IMG;
  g.dy2 = h->dy2; g.z2 = h->a2[1]; g.n2 = Conv2::OUT_PER_IMG;
  g.dy3 = h->dy3; g.z3 = h->a3[1]; g.n3 = Conv3::OUT_PER_IMG;
  g.dh1 = h->dh1; g.z4 = h->h1[1]; g.n4 = kFc1Out;
  g.h1 = h->h1[0]; g.dq_taken = h->dq_taken; g.action = h->act;
  g.v_w2 = v + h->L.fw2; g.v_b2 = v + h->L.fb2; g.out = out;
  adjoint_dot_kernel<<<B, 256, 0, st>>>(g);
  DQNB_CUDA(cudaGetLastError());
  return DQNB_OK;
}

// H . v on the batch of the last update (see hvp_kernels.cuh).  fp32 SIMT GEMMs in both precision modes, one stream.
int enqueue_hvp(dqnb_learner* h, const float* v, float* out, cudaStream_t st) {
  h->ws_off = h->cnt_off = 0;
  const int B = h->B;
  const bool live = !h->sizing;
  dqnb_learner::Hvp& q = h->hvp;
  const float* th = h->online;
  const ArenaLayout& L = h->L;
  auto add_mask = [&](const float* u, const float* w, const float* act, float* o, int64_t n) -> int {
    if (!live) return DQNB_OK;
    hvp_add_mask_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(u, w, act, o, n);
    DQNB_CUDA(cudaGetLastError());
    return DQNB_OK;
  };
  auto conv_lin = [&](auto p, auto xin, const float* w, const float* bias, float* y) -> int {
    using P = decltype(p);
    p.linear = 1;
    p.M = B * P::LAYER_OUT; p.N = P::LAYER_COUT; p.K = P::LAYER_KCOL;
    p.x[0] = xin; p.w[0] = w; p.bias[0] = bias; p.y[0] = y;
    return run_gemm(h, p, 1, pick_splits(gemm_tiles(p, 1), p.K), st);
  };
  auto dense_lin = [&](const float* xin, const float* w, const float* bias, float* y) -> int {
    DenseFwd p;
    p.linear = 1;
    p.M = B; p.N = kFc1Out; p.K = kFc1In;
    p.x[0] = xin; p.w[0] = w; p.bias[0] = bias; p.y[0] = y;
    return run_gemm(h, p, 1, pick_splits(gemm_tiles(p, 1), p.K), st);
  };
  // ---- tangent forward
  DQNB_TRY(conv_lin(HvpConvFwd<Conv1, uint8_t, 64>{}, (const uint8_t*)h->s_tm1, v + L.w1, v + L.b1, q.tu));
  DQNB_TRY(add_mask(q.tu, nullptr, h->a1[0], q.x1d, (int64_t)B * Conv1::OUT_PER_IMG));
  DQNB_TRY(conv_lin(HvpConvFwd<Conv2, float, 32>{}, (const float*)q.x1d, th + L.w2, q.zero, q.tu));
  DQNB_TRY(conv_lin(HvpConvFwd<Conv2, float, 32>{}, (const float*)h->a1[0], v + L.w2, v + L.b2, q.tv));
  DQNB_TRY(add_mask(q.tu, q.tv, h->a2[0], q.x2d, (int64_t)B * Conv2::OUT_PER_IMG));
  DQNB_TRY(conv_lin(HvpConvFwd<Conv3, float, 32>{}, (const float*)q.x2d, th + L.w3, q.zero, q.tu));
  DQNB_TRY(conv_lin(HvpConvFwd<Conv3, float, 32>{}, (const float*)h->a2[0], v + L.w3, v + L.b3, q.tv));
  DQNB_TRY(add_mask(q.tu, q.tv, h->a3[0], q.x3d, (int64_t)B * Conv3::OUT_PER_IMG));
  DQNB_TRY(dense_lin(q.x3d, th + L.fw1, q.zero, q.tu));
  DQNB_TRY(dense_lin(h->a3[0], v + L.fw1, v + L.fb1, q.tv));
  DQNB_TRY(add_mask(q.tu, q.tv, h->h1[0], q.h1d, (int64_t)B * kFc1Out));
  // ---- fc2 / TD head
  if (live) {
    HvpHeadArgs g;
    g.num_actions = h->A; g.shared_bias = h->cfg.shared_bias; g.weight_inside = h->cfg.row_weighted == 2;
    g.grad_bound = h->cfg.grad_error_bound;
    g.inv_mean = 1.0f / (float)(h->cfg.loss_mean_over > 0 ? h->cfg.loss_mean_over : B);
    g.h1 = h->h1[0]; g.h1dot = q.h1d; g.w2 = th + L.fw2; g.v_w2 = v + L.fw2; g.v_b2 = v + L.fb2;
    g.action = h->act; g.td = h->td; g.dq_taken = h->dq_taken; g.row_weight = h->row_weight;
    g.dq_dot = q.dqd; g.dh1_dot = q.dh1d;
    hvp_head_kernel<<<B, 128, 0, st>>>(g);
    const int n = kFc1Out * h->A;
    hvp_fc2_kernel<<<(n + 255) / 256, 256, 0, st>>>(B, h->A, h->cfg.shared_bias, h->h1[0], q.h1d, h->act, h->dq_taken,
                                                    q.dqd, q.ga + L.fw2, q.ga + L.fb2);
    DQNB_CUDA(cudaGetLastError());
  }
  // ---- tangent backward
  auto dense_wgrad = [&](const float* xin, const float* dy, float* grad) -> int {
    DenseWgrad p;
    p.M = kFc1In + 1; p.N = kFc1Out; p.K = B;
    p.x = xin; p.dy = dy; p.grad = grad;
    return run_gemm(h, p, 1, 1, st);
  };
  auto dense_dgrad = [&](const float* dy, const float* w, float* dx) -> int {
    DenseDgrad p;
    p.M = B; p.N = kFc1In; p.K = kFc1Out;
    p.dy = dy; p.w = w; p.act = h->a3[0]; p.dx = dx;
    return run_gemm(h, p, 1, pick_splits(gemm_tiles(p, 1), p.K), st);
  };
  auto conv_wg = [&](auto p, auto xin, const float* dy, float* grad) -> int {
    using P = decltype(p);
    p.M = P::LAYER_KCOL + 1; p.N = P::LAYER_COUT; p.K = B * P::LAYER_OUT;
    p.x = xin; p.dy = dy; p.grad = grad;
    return run_gemm(h, p, 1, pick_splits(gemm_tiles(p, 1), p.K), st);
  };
  auto conv_dg = [&](auto p, const float* dy, const float* w, const float* act_in, float* dx) -> int {
    using P = decltype(p);
    p.M = B * P::NY * P::NX; p.N = P::LAYER_CIN; p.K = P::T * P::T * P::LAYER_COUT;
    p.dy = dy; p.w = w; p.act_in = act_in; p.dx = dx;
    const int Z = P::S * P::S;
    return run_gemm(h, p, Z, pick_splits(gemm_tiles(p, Z), p.K), st);
  };
  DQNB_TRY(dense_wgrad(h->a3[0], q.dh1d, q.ga + L.fw1));
  DQNB_TRY(dense_wgrad(q.x3d, h->dh1, q.gb + L.fw1));
  DQNB_TRY(dense_dgrad(q.dh1d, th + L.fw1, q.tu));
  DQNB_TRY(dense_dgrad(h->dh1, v + L.fw1, q.tv));
  DQNB_TRY(add_mask(q.tu, q.tv, nullptr, q.dy3d, (int64_t)B * Conv3::OUT_PER_IMG));
  DQNB_TRY(conv_wg(HvpConvWgrad<Conv3, float>{}, (const float*)h->a2[0], q.dy3d, q.ga + L.w3));
  DQNB_TRY(conv_wg(HvpConvWgrad<Conv3, float>{}, (const float*)q.x2d, h->dy3, q.gb + L.w3));
  DQNB_TRY(conv_dg(HvpConvDgrad<Conv3>{}, q.dy3d, th + L.w3, h->a2[0], q.tu));
  DQNB_TRY(conv_dg(HvpConvDgrad<Conv3>{}, h->dy3, v + L.w3, h->a2[0], q.tv));
  DQNB_TRY(add_mask(q.tu, q.tv, nullptr, q.dy2d, (int64_t)B * Conv2::OUT_PER_IMG));
  DQNB_TRY(conv_wg(HvpConvWgrad<Conv2, float>{}, (const float*)h->a1[0], q.dy2d, q.ga + L.w2));
  DQNB_TRY(conv_wg(HvpConvWgrad<Conv2, float>{}, (const float*)q.x1d, h->dy2, q.gb + L.w2));
  DQNB_TRY(conv_dg(HvpConvDgrad<Conv2>{}, q.dy2d, th + L.w2, h->a1[0], q.tu));
  DQNB_TRY(conv_dg(HvpConvDgrad<Conv2>{}, h->dy2, v + L.w2, h->a1[0], q.tv));
  DQNB_TRY(add_mask(q.tu, q.tv, nullptr, q.dy1d, (int64_t)B * Conv1::OUT_PER_IMG));
  DQNB_TRY(conv_wg(HvpConvWgrad<Conv1, uint8_t>{}, (const uint8_t*)h->s_tm1, q.dy1d, q.ga + L.w1));
  if (live) {
    HvpBiasRanges r;
    r.lo[0] = L.b1; r.hi[0] = L.w2; r.lo[1] = L.b2; r.hi[1] = L.w3; r.lo[2] = L.b3; r.hi[2] = L.fw1;
    r.lo[3] = L.fb1; r.hi[3] = L.fw2;
    hvp_combine_kernel<<<(unsigned)((L.total + 255) / 256), 256, 0, st>>>(q.ga, q.gb, out, L.total, r);
    DQNB_CUDA(cudaGetLastError());
  }
  return DQNB_OK;
}

int build_graph(dqnb_learner* h, cudaGraphExec_t* out) {
  cudaStream_t cap;
  DQNB_CUDA(cudaStreamCreateWithFlags(&cap, cudaStreamNonBlocking));
  DQNB_CUDA(cudaStreamBeginCapture(cap, cudaStreamCaptureModeThreadLocal));
  int rc = enqueue_any(h, cap);
  cudaGraph_t graph = nullptr;
  cudaError_t e = cudaStreamEndCapture(cap, &graph);
  if (rc != DQNB_OK) { if (graph) cudaGraphDestroy(graph); cudaStreamDestroy(cap); return rc; }
  DQNB_CUDA(e);
  // Per-node priorities: the kernels of the dependent chain outrank the side-branch kernels.
  DQNB_CUDA(cudaGraphInstantiateWithFlags(out, graph, cudaGraphInstantiateFlagUseNodePriority));
  DQNB_CUDA(cudaGraphDestroy(graph));
  DQNB_CUDA(cudaStreamDestroy(cap));
  return DQNB_OK;
}

template <class T>
int dmalloc(T** p, size_t count) {
  DQNB_CUDA(cudaMalloc(p, sizeof(T) * std::max<size_t>(count, 4)));
  DQNB_CUDA(cudaMemset(*p, 0, sizeof(T) * std::max<size_t>(count, 4)));
  return DQNB_OK;
}

}  // namespace

// ------------------------------------------------------------------------- C ABI
extern "C" int dqnb_learner_create(const dqnb_learner_config* cfg, int device, dqnb_learner** out) {
  DQNB_REQUIRE(cfg && out, "null argument");
  DQNB_REQUIRE(cfg->num_actions >= 1 && cfg->num_actions <= kMaxActions, "num_actions %d not in 1..%d",
               cfg->num_actions, kMaxActions);
  DQNB_REQUIRE(cfg->batch_size >= 1 && cfg->batch_size <= 65536, "batch_size %d out of range", cfg->batch_size);
  DQNB_REQUIRE(cfg->precision == DQNB_FP32 || cfg->precision == DQNB_TF32,
               "precision %d not available in this build (DQNB_FP32 or DQNB_TF32)", cfg->precision);
  DeviceGuard guard(device);
  dqnb_learner* h = new dqnb_learner();
  // everything below may fail half way: the handle is destroyed (destroy tolerates partial initialisation)
  const int rc_create = [&]() -> int {
  h->cfg = *cfg;
  h->device = device;
  h->B = cfg->batch_size; h->A = cfg->num_actions; h->nets = cfg->double_q ? 3 : 2;
  h->tf32 = cfg->precision == DQNB_TF32;
  {
    const char* e = getenv("DQNB_STREAM_MIN_BATCH");   // testing knob: batch size from which the ring kernel is used
    h->stream = h->tf32 && cfg->batch_size >= (e ? atoi(e) : kStreamMinBatch);
  }
  h->L = make_layout(h->A, cfg->shared_bias);
  const size_t B = h->B, P = h->L.total;
  DQNB_TRY(dmalloc(&h->online, P)); DQNB_TRY(dmalloc(&h->target, P)); DQNB_TRY(dmalloc(&h->mu, P));
  if (h->tf32) { DQNB_TRY(dmalloc(&h->online_r, P)); DQNB_TRY(dmalloc(&h->target_r, P)); }
  if (h->tf32 && cfg->double_q && getenv("DQNB_NO_F16_CONV1") == nullptr)
    DQNB_TRY(dmalloc(&h->conv1_h16, (size_t)2 * Conv1::COUT * Conv1::KCOL));
  if (h->tf32 && !h->stream && getenv("DQNB_NO_F16_CONV23") == nullptr) {
    DQNB_TRY(dmalloc(&h->conv23_h16, 2 * (dqnb_learner::kC2H + dqnb_learner::kC3H)));
    for (int z = 0; z < h->nets; ++z) {
      DQNB_TRY(dmalloc(&h->a1h[z], B * Conv1::OUT_PER_IMG)); DQNB_TRY(dmalloc(&h->a2h[z], B * Conv2::OUT_PER_IMG));
    }
  }
  if (h->tf32 && !h->stream && h->conv23_h16 != nullptr && getenv("DQNB_NO_TMA") == nullptr &&
      getenv("DQNB_NO_TMA_H16") == nullptr && getenv("DQNB_NO_F16_FC1") == nullptr) {
    DQNB_TRY(dmalloc(&h->fc1_h16, 2 * dqnb_learner::kFc1H));
    for (int z = 0; z < h->nets; ++z) DQNB_TRY(dmalloc(&h->a3h[z], B * Conv3::OUT_PER_IMG));
  }
  if (h->tf32 && !h->stream && getenv("DQNB_NO_TMA") == nullptr) {
    CUtensorMap maps[10] = {};
    DQNB_TRY(make_tmap_2d(h->online_r + h->L.fw1, kFc1Out, kFc1In, 32, 32, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, &maps[0]));
    DQNB_TRY(make_tmap_2d(h->target_r + h->L.fw1, kFc1Out, kFc1In, 32, 32, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, &maps[1]));
    DQNB_TRY(make_tmap_2d(h->online_r + h->L.fw1, kFc1Out, kFc1In, kT32BK, T32DenseDgrad::BN, CU_TENSOR_MAP_SWIZZLE_128B, &maps[2]));
    h->h16_maps = getenv("DQNB_NO_TMA_H16") == nullptr;
    if (h->h16_maps && h->conv1_h16 != nullptr)
      DQNB_TRY(make_tmap_2d(h->conv1_h16, Conv1::KCOL, 2 * Conv1::COUT, 64, 2 * Conv1::COUT, CU_TENSOR_MAP_SWIZZLE_128B, &maps[3]));
    if (h->h16_maps && h->conv23_h16 != nullptr) {
      for (int tgt = 0; tgt < 2; ++tgt) {
        DQNB_TRY(make_tmap_2d(h->c2h(tgt), Conv2::KCOL, Conv2::COUT, 64, Conv2::COUT, CU_TENSOR_MAP_SWIZZLE_128B, &maps[4 + tgt]));
        DQNB_TRY(make_tmap_2d(h->c3h(tgt), Conv3::KCOL, Conv3::COUT, 64, Conv3::COUT, CU_TENSOR_MAP_SWIZZLE_128B, &maps[6 + tgt]));
      }
    }
    if (h->fc1_h16 != nullptr)
      for (int tgt = 0; tgt < 2; ++tgt)
        DQNB_TRY(make_tmap_2d(h->fc1h(tgt), kFc1Out, kFc1In, 64, 64, CU_TENSOR_MAP_SWIZZLE_128B, &maps[8 + tgt]));
    DQNB_CUDA(cudaMalloc(reinterpret_cast<void**>(&h->fc1_maps), sizeof(maps)));
    DQNB_CUDA(cudaMemcpy(h->fc1_maps, maps, sizeof(maps), cudaMemcpyHostToDevice));
  }
  DQNB_TRY(dmalloc(&h->nu, P)); DQNB_TRY(dmalloc(&h->grad, P));
  DQNB_TRY(dmalloc(&h->s_tm1, B * Conv1::IN_PER_IMG)); DQNB_TRY(dmalloc(&h->s_t, B * Conv1::IN_PER_IMG));
  DQNB_TRY(dmalloc(&h->act, B)); DQNB_TRY(dmalloc(&h->rew, B)); DQNB_TRY(dmalloc(&h->disc, B));
  DQNB_TRY(dmalloc(&h->d_slots, B));
  if (cfg->row_weighted) {
    DQNB_TRY(dmalloc(&h->row_weight, B));
    std::vector<float> ones(B, 1.0f);
    DQNB_CUDA(cudaMemcpy(h->row_weight, ones.data(), sizeof(float) * B, cudaMemcpyHostToDevice));
  }
  for (int z = 0; z < h->nets; ++z) {
    DQNB_TRY(dmalloc(&h->a1[z], B * Conv1::OUT_PER_IMG)); DQNB_TRY(dmalloc(&h->a2[z], B * Conv2::OUT_PER_IMG));
    DQNB_TRY(dmalloc(&h->a3[z], B * Conv3::OUT_PER_IMG)); DQNB_TRY(dmalloc(&h->h1[z], B * kFc1Out));
    DQNB_TRY(dmalloc(&h->q[z], B * h->A));
  }
  DQNB_TRY(dmalloc(&h->dh1, B * kFc1Out)); DQNB_TRY(dmalloc(&h->dy3, B * Conv3::OUT_PER_IMG));
  DQNB_TRY(dmalloc(&h->dy2, B * Conv2::OUT_PER_IMG)); DQNB_TRY(dmalloc(&h->dy1, B * Conv1::OUT_PER_IMG));
  DQNB_TRY(dmalloc(&h->td, B)); DQNB_TRY(dmalloc(&h->row_loss, B)); DQNB_TRY(dmalloc(&h->dq_taken, B));
  DQNB_TRY(dmalloc(&h->loss, 1));
  for (auto& s : h->side) DQNB_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  {
    int least = 0, greatest = 0;
    DQNB_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    h->prio_hi = greatest;
    // (numerically lower = scheduled first)  DQNB_BOOST_PRIO = levels below the chain's priority
    const int off = getenv("DQNB_BOOST_PRIO") ? atoi(getenv("DQNB_BOOST_PRIO")) : 0;
    h->prio_boost = std::min(least, greatest + std::max(0, off));
    h->pdl = getenv("DQNB_NO_PDL") == nullptr;
  }
  for (auto& e : h->ev) DQNB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  DQNB_CUDA(cudaEventCreateWithFlags(&h->pinned_free, cudaEventDisableTiming));
  h->pinned_bytes = 2 * B * Conv1::IN_PER_IMG + 16 * B;
  DQNB_CUDA(cudaMallocHost(&h->pinned, h->pinned_bytes));
  DQNB_CUDA(cudaMallocHost(&h->pinned_loss, 64));
  if (cfg->row_weighted) {
    DQNB_CUDA(cudaMallocHost(&h->pinned_w, sizeof(float) * 4 * B));
    for (auto& e : h->w_ev) DQNB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  }
  DQNB_CUDA(cudaHostAlloc(&h->loss_ring, sizeof(unsigned long long) * kLossRing, cudaHostAllocMapped));
  memset(h->loss_ring, 0, sizeof(unsigned long long) * kLossRing);
  DQNB_TRY(dmalloc(&h->d_seq, 1));
  // size the split-K scratch with a dry pass, then record the graph
  h->sizing = true;
  DQNB_TRY(enqueue_any(h, nullptr));
  h->ws_cap = h->ws_off; h->cnt_cap = h->cnt_off; h->part_cap = h->part_off;
  if (cfg->row_weighted == 1) {   // the adjoint-dot GEMMs borrow the same split-K scratch
    const int64_t launches = h->launches_per_update;
    DQNB_TRY(enqueue_adjoint(h, nullptr, nullptr, nullptr));
    h->launches_per_update = launches;
    h->ws_cap = std::max(h->ws_cap, h->ws_off); h->cnt_cap = std::max(h->cnt_cap, h->cnt_off);
  }
  h->sizing = false;
  DQNB_TRY(dmalloc(&h->ws, h->ws_cap)); DQNB_TRY(dmalloc(&h->counters, h->cnt_cap));
  if (h->part_cap) DQNB_TRY(dmalloc(&h->gpart, h->part_cap));
  if (h->fc1_part_cap) DQNB_TRY(dmalloc(&h->fc1_part, h->fc1_part_cap));
  DQNB_CUDA(cudaMalloc(&h->d_tl, sizeof(unsigned long long) * kTlSlot * 64));
  DQNB_TRY(build_graph(h, &h->graph));
  h->tl_graph_n = h->tl_n;
  if (!h->stream && getenv("DQNB_NO_GATHER_OVERLAP") == nullptr) {
    DQNB_TRY(dmalloc(&h->alt_s_tm1, B * Conv1::IN_PER_IMG)); DQNB_TRY(dmalloc(&h->alt_s_t, B * Conv1::IN_PER_IMG));
    DQNB_TRY(dmalloc(&h->alt_act, B)); DQNB_TRY(dmalloc(&h->alt_rew, B)); DQNB_TRY(dmalloc(&h->alt_disc, B));
    h->swap_sets();
    int rc = build_graph(h, &h->graph_alt);
    h->swap_sets();
    DQNB_TRY(rc);
    DQNB_CUDA(cudaStreamCreateWithFlags(&h->gather_st, cudaStreamNonBlocking));
    for (auto& e : h->set_free) DQNB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    DQNB_CUDA(cudaEventCreateWithFlags(&h->gathered, cudaEventDisableTiming));
  }
  return DQNB_OK;
  }();
  if (rc_create != DQNB_OK) { dqnb_learner_destroy(h); return rc_create; }
  *out = h;
  return DQNB_OK;
}

extern "C" int dqnb_learner_destroy(dqnb_learner* h) {
  if (!h) return DQNB_OK;
  DeviceGuard guard(h->device);
  cudaDeviceSynchronize();
  if (h->graph) cudaGraphExecDestroy(h->graph);
  if (h->graph_alt) cudaGraphExecDestroy(h->graph_alt);
  if (h->dp_comm) { const NcclApi* nccl = nccl_api(); if (nccl) nccl->CommDestroy(h->dp_comm); }
  if (h->dp_stream) cudaStreamDestroy(h->dp_stream);
  for (auto& e : h->dp_ev) if (e) cudaEventDestroy(e);
  if (h->act_graph) cudaGraphExecDestroy(h->act_graph);
  if (h->act_obs) cudaFreeHost(h->act_obs);
  if (h->act_q) cudaFreeHost(h->act_q);
  cudaFree(h->act_part); cudaFree(h->act_bar); cudaFree(h->act_seq);
  if (h->act_done) cudaFreeHost(h->act_done);
  if (h->gather_st) cudaStreamDestroy(h->gather_st);
  for (auto& e : h->set_free) if (e) cudaEventDestroy(e);
  if (h->gathered) cudaEventDestroy(h->gathered);
  cudaFree(h->alt_s_tm1); cudaFree(h->alt_s_t); cudaFree(h->alt_act); cudaFree(h->alt_rew); cudaFree(h->alt_disc);
  float* fl[] = {h->online, h->target, h->mu, h->nu, h->grad, h->rew, h->disc, h->dh1, h->dy3, h->dy2, h->dy1,
                 h->td, h->row_loss, h->dq_taken, h->loss, h->ws, h->gpart, h->fc1_part, h->online_r, h->target_r,
                 h->row_weight};
  for (float* p : fl) cudaFree(p);
  for (int z = 0; z < kMaxNets; ++z) { cudaFree(h->a1[z]); cudaFree(h->a2[z]); cudaFree(h->a3[z]); cudaFree(h->h1[z]); cudaFree(h->q[z]); }
  cudaFree(h->s_tm1); cudaFree(h->s_t); cudaFree(h->act); cudaFree(h->counters); cudaFree(h->d_slots);
  {
    dqnb_learner::Hvp& q = h->hvp;
    float* hb[] = {q.x1d, q.x2d, q.x3d, q.h1d, q.dh1d, q.dy3d, q.dy2d, q.dy1d, q.tu, q.tv, q.dqd, q.zero, q.ga, q.gb, q.ws};
    for (float* b : hb) cudaFree(b);
    cudaFree(q.counters);
  }
  cudaFree(h->conv1_h16); cudaFree(h->conv23_h16); cudaFree(h->fc1_maps); cudaFree(h->fc1_h16);
  for (int z = 0; z < kMaxNets; ++z) cudaFree(h->a3h[z]);
  for (int z = 0; z < kMaxNets; ++z) { cudaFree(h->a1h[z]); cudaFree(h->a2h[z]); }
  for (auto& s : h->side) if (s) cudaStreamDestroy(s);
  for (auto& e : h->ev) if (e) cudaEventDestroy(e);
  if (h->pinned_free) cudaEventDestroy(h->pinned_free);
  if (h->pinned) cudaFreeHost(h->pinned);
  if (h->pinned_loss) cudaFreeHost(h->pinned_loss);
  if (h->pinned_w) cudaFreeHost(h->pinned_w);
  for (auto& e : h->w_ev) if (e) cudaEventDestroy(e);
  if (h->loss_ring) cudaFreeHost(h->loss_ring);
  cudaFree(h->d_seq);
  delete h;
  return DQNB_OK;
}

extern "C" int dqnb_learner_get_config(dqnb_learner* h, dqnb_learner_config* cfg, int* device) {
  DQNB_REQUIRE(h && cfg && device, "null argument");
  *cfg = h->cfg;
  *device = h->device;
  return DQNB_OK;
}

extern "C" int dqnb_learner_num_params(dqnb_learner* h, int64_t* n) {
  DQNB_REQUIRE(h && n, "null argument");
  *n = h->L.total;
  return DQNB_OK;
}

namespace {
// fp16 K-major filter copies of one net (fp32 arena `p`) after its weights were written from outside
void refresh_h16(dqnb_learner* h, const float* p, int target, cudaStream_t st) {
  if (h->conv1_h16 != nullptr)
    conv1_h16_kernel<<<(Conv1::KCOL * Conv1::COUT + 255) / 256, 256, 0, st>>>(
        p + h->L.w1, h->conv1_h16 + (target ? Conv1::COUT * Conv1::KCOL : 0), Conv1::KCOL, Conv1::COUT);
  if (h->conv23_h16 != nullptr) {
    conv1_h16_kernel<<<(Conv2::KCOL * Conv2::COUT + 255) / 256, 256, 0, st>>>(p + h->L.w2, h->c2h(target), Conv2::KCOL, Conv2::COUT);
    conv1_h16_kernel<<<(Conv3::KCOL * Conv3::COUT + 255) / 256, 256, 0, st>>>(p + h->L.w3, h->c3h(target), Conv3::KCOL, Conv3::COUT);
  }
  if (h->fc1_h16 != nullptr) {
    const int64_t n = (int64_t)dqnb_learner::kFc1H;
    half_copy_kernel<<<(unsigned)((n / 4 + 255) / 256), 256, 0, st>>>(p + h->L.fw1, h->fc1h(target), n);
  }
}
int locate(dqnb_learner* h, int which, float** p, int64_t* n) {
  const int64_t P = h->L.total, BA = (int64_t)h->B * h->A;
  switch (which) {
    case DQNB_ONLINE: *p = h->online; *n = P; break;
    case DQNB_TARGET: *p = h->target; *n = P; break;
    case DQNB_MU: *p = h->mu; *n = P; break;
    case DQNB_NU: *p = h->nu; *n = P; break;
    case DQNB_GRAD: *p = h->grad; *n = P; break;
    case DQNB_Q_TM1: *p = h->q[0]; *n = BA; break;
    case DQNB_Q_TARGET_T: *p = h->q[1]; *n = BA; break;
    case DQNB_Q_T: DQNB_REQUIRE(h->nets == 3, "Q_T only exists for double-Q learners"); *p = h->q[2]; *n = BA; break;
    case DQNB_TD: *p = h->td; *n = h->B; break;
    case DQNB_LOSS: *p = h->loss; *n = 1; break;
    case DQNB_ROW_WEIGHT:
      DQNB_REQUIRE(h->row_weight, "row weights only exist for learners created with row_weighted = 1");
      *p = h->row_weight; *n = h->B; break;
    // undocumented debug taps (activation gradients / activations of the online net)
    case 100: *p = h->dh1; *n = (int64_t)h->B * kFc1Out; break;
    case 101: *p = h->dy3; *n = (int64_t)h->B * Conv3::OUT_PER_IMG; break;
    case 102: *p = h->dy2; *n = (int64_t)h->B * Conv2::OUT_PER_IMG; break;
    case 103: *p = h->dy1; *n = (int64_t)h->B * Conv1::OUT_PER_IMG; break;
    case 110: *p = h->h1[0]; *n = (int64_t)h->B * kFc1Out; break;
    case 111: *p = h->a3[0]; *n = (int64_t)h->B * Conv3::OUT_PER_IMG; break;
    case 112: *p = h->a2[0]; *n = (int64_t)h->B * Conv2::OUT_PER_IMG; break;
    case 113: *p = h->a1[0]; *n = (int64_t)h->B * Conv1::OUT_PER_IMG; break;
    default: set_error("unknown state array %d", which); return DQNB_EINVAL;
  }
  return DQNB_OK;
}
}  // namespace

extern "C" int dqnb_learner_write(dqnb_learner* h, int which, const float* host, int64_t n, uintptr_t stream) {
  DQNB_REQUIRE(h && host, "null argument");
  DQNB_REQUIRE((which >= DQNB_ONLINE && which <= DQNB_NU) || which == DQNB_ROW_WEIGHT, "state array %d is read-only", which);
  DeviceGuard guard(h->device);
  float* p; int64_t len;
  DQNB_TRY(locate(h, which, &p, &len));
  DQNB_REQUIRE(n == len, "length %lld != %lld", (long long)n, (long long)len);
  DQNB_CUDA(cudaMemcpyAsync(p, host, sizeof(float) * n, cudaMemcpyHostToDevice, as_stream(stream)));
  if (h->tf32 && (which == DQNB_ONLINE || which == DQNB_TARGET)) {
    float* shadow = which == DQNB_ONLINE ? h->online_r : h->target_r;
    round_tf32_kernel<<<(unsigned)((n + 255) / 256), 256, 0, as_stream(stream)>>>(p, shadow, n);
    refresh_h16(h, p, which == DQNB_TARGET, as_stream(stream));
    DQNB_CUDA(cudaGetLastError());
  }
  DQNB_CUDA(cudaStreamSynchronize(as_stream(stream)));
  return DQNB_OK;
}

extern "C" int dqnb_learner_device_ptr(dqnb_learner* h, int which, void** ptr, int64_t* n) {
  DQNB_REQUIRE(h && ptr && n, "null argument");
  float* p; int64_t len;
  DQNB_TRY(locate(h, which, &p, &len));
  *ptr = p; *n = len;
  return DQNB_OK;
}

extern "C" int dqnb_learner_write_dev(dqnb_learner* h, int which, const float* dev_src, int64_t n, uintptr_t stream) {
  DQNB_REQUIRE(h && dev_src, "null argument");
  DQNB_REQUIRE((which >= DQNB_ONLINE && which <= DQNB_GRAD) || which == DQNB_ROW_WEIGHT, "state array %d is read-only", which);
  DeviceGuard guard(h->device);
  float* p; int64_t len;
  DQNB_TRY(locate(h, which, &p, &len));
  DQNB_REQUIRE(n == len, "length %lld != %lld", (long long)n, (long long)len);
  if (p != dev_src) DQNB_CUDA(cudaMemcpyAsync(p, dev_src, sizeof(float) * n, cudaMemcpyDeviceToDevice, as_stream(stream)));
  if (h->tf32 && (which == DQNB_ONLINE || which == DQNB_TARGET)) {
    float* shadow = which == DQNB_ONLINE ? h->online_r : h->target_r;
    round_tf32_kernel<<<(unsigned)((n + 255) / 256), 256, 0, as_stream(stream)>>>(p, shadow, n);
    refresh_h16(h, p, which == DQNB_TARGET, as_stream(stream));
    DQNB_CUDA(cudaGetLastError());
  }
  return DQNB_OK;
}

extern "C" int dqnb_learner_apply_grads(dqnb_learner* h, uintptr_t stream) {
  DQNB_REQUIRE(h, "null argument");
  if (h->dp_on()) {
    set_error("this learner is data-parallel (dqnb_dp_init): its update already all-reduces and applies the gradient");
    return DQNB_ESTATE;
  }
  DeviceGuard guard(h->device);
  cudaStream_t st = as_stream(stream);
  if (h->tf32) {
    GradRegions none{};
    rmsprop_reduce_kernel<true><<<rmsprop_reduce_blocks(h->L.total, 0, 0, 0), 256, 0, st>>>(
        h->online, h->online_r, h->h16_online(), h->mu, h->nu, h->grad, h->L.total, none, 0, 0, 0, h->cfg.learning_rate,
        h->cfg.decay, h->cfg.eps, nullptr);
  } else {
    int64_t n4 = (h->L.total >> 2) + 1;
    rmsprop_kernel<<<(unsigned)((n4 + 255) / 256), 256, 0, st>>>(h->online, h->mu, h->nu, h->grad, h->L.total,
                                                               h->cfg.learning_rate, h->cfg.decay, h->cfg.eps, nullptr);
  }
  DQNB_CUDA(cudaGetLastError());
  return DQNB_OK;
}

// ------------------------------------------------------------------ data-parallel learner
extern "C" int dqnb_dp_unique_id(uint8_t* id128) {
  DQNB_REQUIRE(id128, "null argument");
  const NcclApi* nccl = nccl_api();
  if (!nccl) return DQNB_ESTATE;
  NcclUniqueId id;
  DQNB_NCCL(nccl, nccl->GetUniqueId(&id));
  memcpy(id128, id.internal, sizeof(id.internal));
  return DQNB_OK;
}

extern "C" int dqnb_dp_init(dqnb_learner* h, const uint8_t* id128, int rank, int world) {
  DQNB_REQUIRE(h && id128, "null argument");
  DQNB_REQUIRE(world >= 1 && rank >= 0 && rank < world, "rank %d not in a world of %d", rank, world);
  DQNB_REQUIRE(h->cfg.defer_apply, "create the learner with defer_apply = 1 (and loss_mean_over = the global batch)");
  DQNB_REQUIRE(!h->dp_on(), "dqnb_dp_init was already called on this learner");
  const NcclApi* nccl = nccl_api();
  if (!nccl) return DQNB_ESTATE;
  DeviceGuard guard(h->device);
  DQNB_CUDA(cudaDeviceSynchronize());
  NcclUniqueId id;
  memcpy(id.internal, id128, sizeof(id.internal));
  NcclComm comm = nullptr;
  DQNB_NCCL(nccl, nccl->CommInitRank(&comm, world, id, rank));
  h->dp_comm = comm; h->dp_rank = rank; h->dp_world = world;
  DQNB_CUDA(cudaStreamCreateWithPriority(&h->dp_stream, cudaStreamNonBlocking, h->prio_hi));
  for (auto& e : h->dp_ev) DQNB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  // one eager collective first: NCCL sets up its channels / buffers on the first call, which must not happen
  // inside a stream capture
  DQNB_NCCL(nccl, nccl->AllReduce(h->grad, h->grad, (size_t)h->L.total, kNcclFloat32, kNcclSum, h->dp_comm, h->dp_stream));
  DQNB_CUDA(cudaStreamSynchronize(h->dp_stream));
  DQNB_CUDA(cudaMemsetAsync(h->grad, 0, sizeof(float) * h->L.total, h->dp_stream));
  DQNB_CUDA(cudaStreamSynchronize(h->dp_stream));
  // re-record the update graph(s): backward -> [reduce + all-reduce] x 2 -> optimiser step
  if (h->graph) { cudaGraphExecDestroy(h->graph); h->graph = nullptr; }
  DQNB_TRY(build_graph(h, &h->graph));
  h->tl_graph_n = h->tl_n;
  if (h->graph_alt) {
    cudaGraphExecDestroy(h->graph_alt); h->graph_alt = nullptr;
    h->swap_sets();
    int rc = build_graph(h, &h->graph_alt);
    h->swap_sets();
    DQNB_TRY(rc);
  }
  return DQNB_OK;
}

// The collectives of one data-parallel update alone (same two chunks, same stream), for timing.
extern "C" int dqnb_dp_allreduce(dqnb_learner* h, uintptr_t stream) {
  DQNB_REQUIRE(h && h->dp_on(), "not a data-parallel learner (dqnb_dp_init)");
  const NcclApi* nccl = nccl_api();
  if (!nccl) return DQNB_ESTATE;
  DeviceGuard guard(h->device);
  cudaStream_t st = as_stream(stream);
  DQNB_NCCL(nccl, nccl->AllReduce(h->grad + h->L.fw1, h->grad + h->L.fw1, (size_t)(h->L.total - h->L.fw1),
                                  kNcclFloat32, kNcclSum, h->dp_comm, st));
  DQNB_NCCL(nccl, nccl->AllReduce(h->grad, h->grad, (size_t)h->L.fw1, kNcclFloat32, kNcclSum, h->dp_comm, st));
  return DQNB_OK;
}

extern "C" int dqnb_dp_info(dqnb_learner* h, int* rank, int* world, int* nccl_version) {
  DQNB_REQUIRE(h && rank && world && nccl_version, "null argument");
  *rank = h->dp_rank; *world = h->dp_world; *nccl_version = 0;
  if (h->dp_on()) {
    const NcclApi* nccl = nccl_api();
    if (nccl) nccl->GetVersion(nccl_version);
  }
  return DQNB_OK;
}

// H . v, H = Jacobian of the loss gradient w.r.t. the online parameters at the batch (and td errors, clip decisions,
// ReLU masks) of the LAST update of this learner; v_dev / out_dev are device vectors laid out like the parameters.
extern "C" int dqnb_learner_hvp(dqnb_learner* h, const float* v_dev, float* out_dev, uintptr_t stream) {
  DQNB_REQUIRE(h && v_dev && out_dev, "null argument");
  DQNB_REQUIRE(h->updates > 0, "the Hessian-vector product is taken at the batch of the last update: run one first");
  DeviceGuard guard(h->device);
  dqnb_learner::Hvp& q = h->hvp;
  float* const ws_keep = h->ws; unsigned* const cnt_keep = h->counters;
  const int64_t launches = h->launches_per_update;
  if (!q.ready) {
    const size_t B = h->B, P = h->L.total;
    const size_t big = B * Conv1::OUT_PER_IMG;
    DQNB_TRY(dmalloc(&q.x1d, B * Conv1::OUT_PER_IMG)); DQNB_TRY(dmalloc(&q.x2d, B * Conv2::OUT_PER_IMG));
    DQNB_TRY(dmalloc(&q.x3d, B * Conv3::OUT_PER_IMG)); DQNB_TRY(dmalloc(&q.h1d, B * kFc1Out));
    DQNB_TRY(dmalloc(&q.dh1d, B * kFc1Out)); DQNB_TRY(dmalloc(&q.dy3d, B * Conv3::OUT_PER_IMG));
    DQNB_TRY(dmalloc(&q.dy2d, B * Conv2::OUT_PER_IMG)); DQNB_TRY(dmalloc(&q.dy1d, B * Conv1::OUT_PER_IMG));
    DQNB_TRY(dmalloc(&q.tu, big)); DQNB_TRY(dmalloc(&q.tv, big)); DQNB_TRY(dmalloc(&q.dqd, B));
    DQNB_TRY(dmalloc(&q.zero, (size_t)kFc1Out)); DQNB_TRY(dmalloc(&q.ga, P)); DQNB_TRY(dmalloc(&q.gb, P));
    h->sizing = true;
    int rc = enqueue_hvp(h, v_dev, out_dev, nullptr);
    h->sizing = false;
    h->launches_per_update = launches;
    DQNB_TRY(rc);
    DQNB_TRY(dmalloc(&q.ws, h->ws_off)); DQNB_TRY(dmalloc(&q.counters, h->cnt_off));
    q.ready = true;
  }
  h->ws = q.ws; h->counters = q.counters;
  const int rc = enqueue_hvp(h, v_dev, out_dev, as_stream(stream));
  h->ws = ws_keep; h->counters = cnt_keep;
  h->launches_per_update = launches;
  return rc;
}

extern "C" int dqnb_learner_adjoint_dots(dqnb_learner* h, const float* v_dev, float* out_dev, uintptr_t stream) {
  DQNB_REQUIRE(h && v_dev && out_dev, "null argument");
  DQNB_REQUIRE(h->row_weight, "adjoint dots need a learner created with row_weighted = 1");
  DQNB_REQUIRE(h->nets == 2, "adjoint dots are defined for the q_learning loss (double_q = 0)");
  DeviceGuard guard(h->device);
  return enqueue_adjoint(h, v_dev, out_dev, as_stream(stream));
}

extern "C" int dqnb_learner_read(dqnb_learner* h, int which, float* host, int64_t n, uintptr_t stream) {
  DQNB_REQUIRE(h && host, "null argument");
  DeviceGuard guard(h->device);
  float* p; int64_t len;
  DQNB_TRY(locate(h, which, &p, &len));
  DQNB_REQUIRE(n == len, "length %lld != %lld", (long long)n, (long long)len);
  DQNB_CUDA(cudaMemcpyAsync(host, p, sizeof(float) * n, cudaMemcpyDeviceToHost, as_stream(stream)));
  DQNB_CUDA(cudaStreamSynchronize(as_stream(stream)));
  return DQNB_OK;
}

extern "C" int dqnb_learner_update(dqnb_learner* h, const uint8_t* s_tm1_dev, const int32_t* a_dev,
                                   const float* r_dev, const float* disc_dev, const uint8_t* s_t_dev,
                                   uintptr_t stream) {
  DQNB_REQUIRE(h && s_tm1_dev && a_dev && r_dev && disc_dev && s_t_dev, "null argument");
  DeviceGuard guard(h->device);
  cudaStream_t st = as_stream(stream);
  const size_t B = h->B, img = (size_t)Conv1::IN_PER_IMG * B;
  if (s_tm1_dev != h->s_tm1) {   // external batch: bring it into the graph's fixed buffers
    DQNB_CUDA(cudaMemcpyAsync(h->s_tm1, s_tm1_dev, img, cudaMemcpyDeviceToDevice, st));
    DQNB_CUDA(cudaMemcpyAsync(h->s_t, s_t_dev, img, cudaMemcpyDeviceToDevice, st));
    DQNB_CUDA(cudaMemcpyAsync(h->act, a_dev, 4 * B, cudaMemcpyDeviceToDevice, st));
    DQNB_CUDA(cudaMemcpyAsync(h->rew, r_dev, 4 * B, cudaMemcpyDeviceToDevice, st));
    DQNB_CUDA(cudaMemcpyAsync(h->disc, disc_dev, 4 * B, cudaMemcpyDeviceToDevice, st));
  }
  DQNB_CUDA(cudaGraphLaunch(h->graph, st));
  if (h->graph_alt) { DQNB_CUDA(cudaEventRecord(h->set_free[0], st)); h->set_used[0] = true; }
  h->updates++;
  h->seq_host++;
  return DQNB_OK;
}

extern "C" int dqnb_learner_update_host(dqnb_learner* h, const uint8_t* s_tm1_host, const int32_t* a_host,
                                        const float* r_host, const float* disc_host, const uint8_t* s_t_host,
                                        uintptr_t stream) {
  DQNB_REQUIRE(h && s_tm1_host && a_host && r_host && disc_host && s_t_host, "null argument");
  DeviceGuard guard(h->device);
  cudaStream_t st = as_stream(stream);
  const size_t B = h->B, img = (size_t)Conv1::IN_PER_IMG * B;
  if (h->pinned_busy) DQNB_CUDA(cudaEventSynchronize(h->pinned_free));
  uint8_t* p = h->pinned;
  memcpy(p, s_tm1_host, img); memcpy(p + img, s_t_host, img);
  memcpy(p + 2 * img, a_host, 4 * B); memcpy(p + 2 * img + 4 * B, r_host, 4 * B);
  memcpy(p + 2 * img + 8 * B, disc_host, 4 * B);
  DQNB_CUDA(cudaMemcpyAsync(h->s_tm1, p, img, cudaMemcpyHostToDevice, st));
  DQNB_CUDA(cudaMemcpyAsync(h->s_t, p + img, img, cudaMemcpyHostToDevice, st));
  DQNB_CUDA(cudaMemcpyAsync(h->act, p + 2 * img, 4 * B, cudaMemcpyHostToDevice, st));
  DQNB_CUDA(cudaMemcpyAsync(h->rew, p + 2 * img + 4 * B, 4 * B, cudaMemcpyHostToDevice, st));
  DQNB_CUDA(cudaMemcpyAsync(h->disc, p + 2 * img + 8 * B, 4 * B, cudaMemcpyHostToDevice, st));
  DQNB_CUDA(cudaEventRecord(h->pinned_free, st));
  h->pinned_busy = true;
  return dqnb_learner_update(h, h->s_tm1, h->act, h->rew, h->disc, h->s_t, stream);
}

extern "C" int dqnb_learner_update_from_replay(dqnb_learner* h, dqnb_replay* replay, const int64_t* slots_host,
                                               uintptr_t stream) {
  DQNB_REQUIRE(h && replay && slots_host, "null argument");
  int H, W, S;
  replay_frame_shape(replay, &H, &W, &S);
  DQNB_REQUIRE(H == Conv1::HIN && W == Conv1::WIN && S == Conv1::CIN, "replay frames %dx%dx%d are not 84x84x4", H, W, S);
  DeviceGuard guard(h->device);
  if (h->graph_alt == nullptr) {
    DQNB_TRY(replay_gather_into(replay, slots_host, h->B, h->s_tm1, h->s_t, h->act, h->rew, h->disc, as_stream(stream)));
    return dqnb_learner_update(h, h->s_tm1, h->act, h->rew, h->disc, h->s_t, stream);
  }
  // Gather into the buffer set the running update is NOT reading, on the gather stream: it waits only
  // for the last reader of that set (two updates ago) and for the replay's own commits.
  cudaStream_t st = as_stream(stream), gs = h->gather_st;
  const int k = h->flip;
  if (h->set_used[k]) DQNB_CUDA(cudaStreamWaitEvent(gs, h->set_free[k], 0));
  if (k) DQNB_TRY(replay_gather_into(replay, slots_host, h->B, h->alt_s_tm1, h->alt_s_t, h->alt_act, h->alt_rew, h->alt_disc, gs));
  else DQNB_TRY(replay_gather_into(replay, slots_host, h->B, h->s_tm1, h->s_t, h->act, h->rew, h->disc, gs));
  DQNB_CUDA(cudaEventRecord(h->gathered, gs));
  DQNB_CUDA(cudaStreamWaitEvent(st, h->gathered, 0));
  DQNB_CUDA(cudaGraphLaunch(k ? h->graph_alt : h->graph, st));
  DQNB_CUDA(cudaEventRecord(h->set_free[k], st));
  h->set_used[k] = true;
  h->flip ^= 1;
  h->updates++;
  h->seq_host++;
  return DQNB_OK;
}

// `_learn` of the prioritized agent (prioritized/agent.py:215-236): as update_from_replay, with this batch's
// importance-sampling weights (host float32 [batch]) staged into DQNB_ROW_WEIGHT ahead of the step.
extern "C" int dqnb_learner_update_from_replay_weighted(dqnb_learner* h, dqnb_replay* replay, const int64_t* slots_host,
                                                        const float* weights_host, uintptr_t stream) {
  DQNB_REQUIRE(h && weights_host, "null argument");
  DQNB_REQUIRE(h->row_weight && h->pinned_w, "row weights need a learner created with row_weighted = 1 or 2");
  DeviceGuard guard(h->device);
  cudaStream_t st = as_stream(stream);
  const int k = h->w_cur;
  h->w_cur = (k + 1) & 3;
  if (h->w_busy[k]) DQNB_CUDA(cudaEventSynchronize(h->w_ev[k]));
  float* pw = h->pinned_w + (size_t)k * h->B;
  memcpy(pw, weights_host, sizeof(float) * h->B);
  DQNB_CUDA(cudaMemcpyAsync(h->row_weight, pw, sizeof(float) * h->B, cudaMemcpyHostToDevice, st));
  DQNB_CUDA(cudaEventRecord(h->w_ev[k], st));
  h->w_busy[k] = true;
  return dqnb_learner_update_from_replay(h, replay, slots_host, stream);
}

extern "C" int dqnb_learner_loss_post(dqnb_learner* h, uintptr_t) {
  DQNB_REQUIRE(h, "null argument");
  DQNB_REQUIRE(h->seq_host > 0, "no update launched yet");
  DQNB_REQUIRE(h->loss_posted - h->loss_taken < kLossRing, "%d loss read-backs already outstanding: take one first",
               kLossRing);
  h->loss_want[h->loss_posted & (kLossRing - 1)] = h->seq_host;
  h->loss_posted++;
  return DQNB_OK;
}

extern "C" int dqnb_learner_loss_take(dqnb_learner* h, float* loss) {
  DQNB_REQUIRE(h && loss, "null argument");
  DQNB_REQUIRE(h->loss_posted > h->loss_taken, "no loss read-back outstanding");
  const uint32_t want = h->loss_want[h->loss_taken & (kLossRing - 1)];
  h->loss_taken++;
  volatile unsigned long long* slot = h->loss_ring + (want & (kLossRing - 1));
  const auto t0 = std::chrono::steady_clock::now();
  for (uint64_t spins = 0;; ++spins) {
    const unsigned long long v = *slot;
    const uint32_t seq = (uint32_t)(v >> 32);
    if (seq == want) {
      const uint32_t bits = (uint32_t)v;
      memcpy(loss, &bits, sizeof(float));
      return DQNB_OK;
    }
    DQNB_REQUIRE((int32_t)(seq - want) < 0, "loss read-back overrun: %d more updates were launched before the take",
                 kLossRing);
    if ((spins & 1023) == 1023) {
      DeviceGuard guard(h->device);
      DQNB_CUDA(cudaGetLastError());   // surface a faulted device instead of spinning forever
      DQNB_REQUIRE(std::chrono::steady_clock::now() - t0 < std::chrono::seconds(30), "loss read-back timed out");
    }
  }
}

extern "C" int dqnb_learner_sync_target(dqnb_learner* h, uintptr_t stream) {
  DQNB_REQUIRE(h, "null argument");
  DeviceGuard guard(h->device);
  DQNB_CUDA(cudaMemcpyAsync(h->target, h->online, sizeof(float) * h->L.total, cudaMemcpyDeviceToDevice, as_stream(stream)));
  if (h->tf32)
    DQNB_CUDA(cudaMemcpyAsync(h->target_r, h->online_r, sizeof(float) * h->L.total, cudaMemcpyDeviceToDevice,
                              as_stream(stream)));
  if (h->conv1_h16 != nullptr)
    DQNB_CUDA(cudaMemcpyAsync(h->conv1_h16 + Conv1::COUT * Conv1::KCOL, h->conv1_h16,
                              sizeof(__half) * Conv1::COUT * Conv1::KCOL, cudaMemcpyDeviceToDevice, as_stream(stream)));
  if (h->conv23_h16 != nullptr)
    DQNB_CUDA(cudaMemcpyAsync(h->c2h(1), h->c2h(0), sizeof(__half) * (dqnb_learner::kC2H + dqnb_learner::kC3H),
                              cudaMemcpyDeviceToDevice, as_stream(stream)));
  if (h->fc1_h16 != nullptr)
    DQNB_CUDA(cudaMemcpyAsync(h->fc1h(1), h->fc1h(0), sizeof(__half) * dqnb_learner::kFc1H, cudaMemcpyDeviceToDevice,
                              as_stream(stream)));
  return DQNB_OK;
}

namespace {
// The forward of `n` observations already in h->s_tm1 through one net, q into h->q[0] (net-0 buffers
// are borrowed: acting runs between updates on the same stream).
int enqueue_q_forward(dqnb_learner* h, int which_params, int n, cudaStream_t st) {
  const uint8_t* xin[kMaxNets] = {h->s_tm1, h->s_tm1, h->s_tm1};
  const float* prm = which_params ? h->target : h->online;
  const float* params[kMaxNets] = {prm, prm, prm};
  h->ws_off = h->cnt_off = 0;
  const float* prm_r = which_params ? h->target_r : h->online_r;
  const float* wts[kMaxNets] = {prm_r, prm_r, prm_r};
  if (h->tf32) DQNB_TRY(forward_torso_tf32(h, 1, n, xin, params, wts, st));
  else DQNB_TRY(forward_torso(h, 1, n, xin, params, st));
  HeadArgs g = head_args(h, n, params, 2);
  g.h1[1] = g.h1[0]; g.q[1] = h->q[1];   // head also evaluates a dummy target row; harmless
  g.row_weight = nullptr;
  if (h->tf32) {   // one net's partials; the dummy second net re-reads them
    g.fc1_part = h->fc1_part; g.fc1_splits = h->fc1_splits;
  }
  g.td = h->td; g.dh1 = h->dh1;
  head_td_for(n)<<<n, kHeadThreads, 0, st>>>(g);
  DQNB_CUDA(cudaGetLastError());
  return DQNB_OK;
}

int build_act_graph(dqnb_learner* h) {
  if (h->act_obs == nullptr) DQNB_CUDA(cudaMallocHost(&h->act_obs, Conv1::IN_PER_IMG));
  if (h->act_q == nullptr) DQNB_CUDA(cudaHostAlloc(&h->act_q, sizeof(float) * kMaxActions, cudaHostAllocMapped));
  if (h->act_part == nullptr) DQNB_TRY(dmalloc(&h->act_part, (size_t)kActSlices * kFc1Out));
  if (h->act_bar == nullptr) DQNB_TRY(dmalloc(&h->act_bar, 1));
  if (h->act_seq == nullptr) DQNB_TRY(dmalloc(&h->act_seq, 1));
  if (h->act_done == nullptr) {
    DQNB_CUDA(cudaHostAlloc(&h->act_done, sizeof(unsigned), cudaHostAllocMapped));
    *h->act_done = 0;
  }
  int sms = 0, per_sm = 0;
  DQNB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device));
  DQNB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, act_forward_kernel, kActThreads, 0));
  DQNB_REQUIRE(per_sm >= 1 && sms >= 1, "acting kernel does not fit on an SM");
  ActArgs a;
  a.obs = h->s_tm1;
  a.w1 = h->online + h->L.w1; a.b1 = h->online + h->L.b1; a.w2 = h->online + h->L.w2; a.b2 = h->online + h->L.b2;
  a.w3 = h->online + h->L.w3; a.b3 = h->online + h->L.b3; a.fw1 = h->online + h->L.fw1; a.fb1 = h->online + h->L.fb1;
  a.fw2 = h->online + h->L.fw2; a.fb2 = h->online + h->L.fb2;
  a.a1 = h->a1[0]; a.a2 = h->a2[0]; a.a3 = h->a3[0]; a.part = h->act_part; a.q = h->q[0];
  a.q_host = h->act_q;   // unified addressing: the pinned allocation is device-accessible at the same address
  a.bar = h->act_bar; a.num_actions = h->A; a.shared_bias = h->cfg.shared_bias;
  a.seq_dev = h->act_seq; a.done_host = h->act_done;
  const int grid = std::min(sms, kNumSMs);
  cudaStream_t cap;
  DQNB_CUDA(cudaStreamCreateWithFlags(&cap, cudaStreamNonBlocking));
  DQNB_CUDA(cudaStreamBeginCapture(cap, cudaStreamCaptureModeThreadLocal));
  h->chain = cap;   // consecutive kernels: chain priority + programmatic dependent launch
  int rc = DQNB_OK;
  if (cudaMemcpyAsync(h->s_tm1, h->act_obs, Conv1::IN_PER_IMG, cudaMemcpyHostToDevice, cap) != cudaSuccess) rc = DQNB_ECUDA;
  if (rc == DQNB_OK) {
    // cooperative launch: the grid barrier needs every CTA resident at once, and only a cooperative grid is
    // scheduled all-or-nothing (two plain launches from different streams could each hold half the SMs)
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3((unsigned)grid); lc.blockDim = dim3(kActThreads); lc.stream = cap;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
    lc.attrs = attr; lc.numAttrs = 1;
    if (cudaLaunchKernelEx(&lc, act_forward_kernel, a) != cudaSuccess) { cudaGetLastError(); rc = DQNB_ECUDA; set_error("cooperative launch of the acting kernel failed"); }
  }
  cudaGraph_t graph = nullptr;
  cudaError_t e = cudaStreamEndCapture(cap, &graph);
  h->chain = nullptr;
  if (rc != DQNB_OK || e != cudaSuccess) {
    if (graph) cudaGraphDestroy(graph);
    cudaStreamDestroy(cap);
    cudaGetLastError();
    if (rc == DQNB_OK) { set_error("capture of the acting graph failed: %s", cudaGetErrorString(e)); rc = DQNB_ECUDA; }
    return rc;
  }
  DQNB_CUDA(cudaGraphInstantiateWithFlags(&h->act_graph, graph, cudaGraphInstantiateFlagUseNodePriority));
  DQNB_CUDA(cudaGraphDestroy(graph));
  DQNB_CUDA(cudaStreamDestroy(cap));
  return DQNB_OK;
}
}  // namespace

extern "C" int dqnb_learner_q_values(dqnb_learner* h, int which_params, const uint8_t* s_host, int32_t n,
                                     float* q_host, uintptr_t stream) {
  DQNB_REQUIRE(h && s_host && q_host, "null argument");
  DQNB_REQUIRE(n >= 1 && n <= h->B, "n=%d not in 1..batch_size", n);
  DQNB_REQUIRE(which_params == 0 || which_params == 1, "which_params must be 0 (online) or 1 (target)");
  DeviceGuard guard(h->device);
  cudaStream_t st = as_stream(stream);
  if (n == 1 && which_params == 0 && getenv("DQNB_NO_ACT_GRAPH") == nullptr) {
    if (h->act_graph == nullptr) DQNB_TRY(build_act_graph(h));
    memcpy(h->act_obs, s_host, Conv1::IN_PER_IMG);
    DQNB_CUDA(cudaGraphLaunch(h->act_graph, st));
    if (h->graph_alt) { DQNB_CUDA(cudaEventRecord(h->set_free[0], st)); h->set_used[0] = true; }
    // poll the kernel's completion flag in pinned memory (a stream synchronise costs several us of wake-up)
    const unsigned want = ++h->act_launched;
    const volatile unsigned* done = h->act_done;
    const auto t0 = std::chrono::steady_clock::now();
    for (uint64_t spins = 0; *done != want; ++spins) {
      if ((spins & 4095) == 4095) {
        if (cudaStreamQuery(st) != cudaErrorNotReady) break;   // finished (flag lands with it) or faulted
        DQNB_REQUIRE(std::chrono::steady_clock::now() - t0 < std::chrono::seconds(30), "acting kernel timed out");
      }
    }
    if (*done != want) DQNB_CUDA(cudaStreamSynchronize(st));
    DQNB_REQUIRE(*done == want, "acting kernel finished without reporting (sequence %u, expected %u)", *done, want);
    memcpy(q_host, h->act_q, sizeof(float) * h->A);
    return DQNB_OK;
  }
  // acting runs between updates on the same stream, so net-0 buffers can be borrowed
  DQNB_CUDA(cudaMemcpyAsync(h->s_tm1, s_host, (size_t)n * Conv1::IN_PER_IMG, cudaMemcpyHostToDevice, st));
  DQNB_TRY(enqueue_q_forward(h, which_params, n, st));
  if (h->graph_alt) { DQNB_CUDA(cudaEventRecord(h->set_free[0], st)); h->set_used[0] = true; }
  DQNB_CUDA(cudaMemcpyAsync(q_host, h->q[0], sizeof(float) * n * h->A, cudaMemcpyDeviceToHost, st));
  DQNB_CUDA(cudaStreamSynchronize(st));
  return DQNB_OK;
}

extern "C" int dqnb_learner_relaunch(dqnb_learner* h, int times, uintptr_t stream) {
  DQNB_REQUIRE(h && times >= 0, "bad argument");
  DeviceGuard guard(h->device);
  for (int i = 0; i < times; ++i) {
    DQNB_CUDA(cudaGraphLaunch(h->graph, as_stream(stream)));
    h->updates++;
    h->seq_host++;
  }
  if (h->graph_alt && times > 0) { DQNB_CUDA(cudaEventRecord(h->set_free[0], as_stream(stream))); h->set_used[0] = true; }
  return DQNB_OK;
}

extern "C" int dqnb_learner_launches_per_update(dqnb_learner* h, int64_t* n) {
  DQNB_REQUIRE(h && n, "null argument");
  *n = h->launches_per_update;
  return DQNB_OK;
}

// Per-kernel timing of one update: the same launches as the graph, but eagerly on one stream
// with a CUDA event after each, averaged over `iters` runs (the first is a warm-up).  State is
// advanced like by `iters` updates on the batch currently in the learner's buffers.
extern "C" int dqnb_learner_profile(dqnb_learner* h, int iters, char* report, size_t report_len, uintptr_t stream) {
  DQNB_REQUIRE(h && report && report_len > 0 && iters >= 1, "bad argument");
  DeviceGuard guard(h->device);
  cudaStream_t st = as_stream(stream);
  cudaEvent_t begin;
  DQNB_CUDA(cudaEventCreate(&begin));
  std::vector<double> total;
  if (!h->dbg_clocks) DQNB_CUDA(cudaMalloc(&h->dbg_clocks, sizeof(long long) * 8 * 64));
  DQNB_CUDA(cudaMemset(h->dbg_clocks, 0, sizeof(long long) * 8 * 64));
  h->profiling = true;
  int rc = DQNB_OK;
  for (int it = 0; it <= iters && rc == DQNB_OK; ++it) {
    h->prof_n = 0;
    cudaEventRecord(begin, st);
    rc = enqueue_any(h, st);
    if (rc != DQNB_OK) break;
    h->seq_host++;
    if (cudaStreamSynchronize(st) != cudaSuccess) { set_error("profile: stream sync failed"); rc = DQNB_ECUDA; break; }
    if (it == 0) { total.assign(h->prof_n, 0.0); continue; }
    for (size_t i = 0; i < h->prof_n; ++i) {
      float ms = 0.f;
      cudaEventElapsedTime(&ms, i == 0 ? begin : h->prof_ev[i - 1], h->prof_ev[i]);
      total[i] += ms * 1e3;
    }
  }
  h->profiling = false;
  cudaEventDestroy(begin);
  DQNB_TRY(rc);
  size_t off = 0;
  double sum = 0;
  std::vector<long long> clk(8 * 64, 0);
  cudaMemcpy(clk.data(), h->dbg_clocks, sizeof(long long) * 8 * 64, cudaMemcpyDeviceToHost);
  for (size_t i = 0; i < h->prof_n && off + 200 < report_len; ++i) {
    off += snprintf(report + off, report_len - off, "%-16s %8.2f us", h->prof_name[i], total[i] / iters);
    const long long* c = &clk[8 * i];
    if (i < 64 && c[0] != 0)   // CTA 0 phase durations in SM cycles (last iteration)
      off += snprintf(report + off, report_len - off,
                      "   cta0 cycles: setup %lld issue %lld land %lld mma %lld epi %lld fixup %lld (sync1 %lld)",
                      c[1] - c[0], c[2] - c[1], c[3] - c[2], c[4] - c[3], c[5] - c[4], c[6] - c[5],
                      c[7] ? c[7] - c[5] : 0LL);
    off += snprintf(report + off, report_len - off, "\n");
    sum += total[i] / iters;
  }
  if (off + 64 < report_len) snprintf(report + off, report_len - off, "%-16s %8.2f us\n", "sum", sum);
  return DQNB_OK;
}

// Timeline of one *graph* update as it really runs (parallel branches, launch gaps): begin / end
// of every kernel in microseconds relative to the first kernel's begin, from %globaltimer stamps.
extern "C" int dqnb_learner_timeline(dqnb_learner* h, char* report, size_t report_len, uintptr_t stream) {
  DQNB_REQUIRE(h && report && report_len > 0, "bad argument");
  DeviceGuard guard(h->device);
  cudaStream_t st = as_stream(stream);
  const int n = h->tl_graph_n;
  std::vector<unsigned long long> init(kTlSlot * 64, 0ull), got(kTlSlot * 64);
  for (int i = 0; i < 64; ++i) init[kTlSlot * i] = ~0ull;
  const size_t bytes = sizeof(unsigned long long) * kTlSlot * 64;
  for (int rep = 0; rep < 3; ++rep) {   // the last repetition is reported
    DQNB_CUDA(cudaMemcpyAsync(h->d_tl, init.data(), bytes, cudaMemcpyHostToDevice, st));
    DQNB_CUDA(cudaStreamSynchronize(st));
    DQNB_CUDA(cudaGraphLaunch(h->graph, st));
    h->seq_host++;
    DQNB_CUDA(cudaStreamSynchronize(st));
  }
  DQNB_CUDA(cudaMemcpy(got.data(), h->d_tl, bytes, cudaMemcpyDeviceToHost));
  unsigned long long t0 = ~0ull, t1 = 0;
  for (int i = 0; i < n; ++i) { t0 = std::min(t0, got[kTlSlot * i]); t1 = std::max(t1, got[kTlSlot * i + 1]); }
  size_t off = 0;
  for (int i = 0; i < n && off + 128 < report_len; ++i) {
    const unsigned long long* g = &got[kTlSlot * i];
    off += snprintf(report + off, report_len - off,
                    "%-16s begin %8.2f  end %8.2f  (%6.2f us)  last CTA starts +%5.2f, longest CTA %5.2f\n", h->tl_name[i],
                    (g[0] - t0) * 1e-3, (g[1] - t0) * 1e-3, (g[1] - g[0]) * 1e-3, (g[2] - g[0]) * 1e-3, g[3] * 1e-3);
  }
  if (off + 64 < report_len) snprintf(report + off, report_len - off, "span %.2f us\n", (t1 - t0) * 1e-3);
  return DQNB_OK;
}
