/* dqnb200.h — C ABI of libdqnb200.so: the B200 (sm_100a) DQN learner hot path.
 *
 * The reference (Stalfoes/dqn_mgsc_zoo) is pure Python with no FFI; its seams for
 * this path are duck-typed Python objects.  Each entry point below names the
 * reference interface (file:line under /root/reference/dqn_zoo) it stands in for.
 * INTEGRATION.md shows the ctypes binding a maintainer would add.
 *
 * Conventions: every function returns 0 on success or a negative DQNB_E* code;
 * dqnb_last_error() returns a thread-local message for the last failure.  No C++
 * types or exceptions cross the boundary.  Handles are opaque and not
 * thread-safe per handle.  `stream` is a cudaStream_t passed as uintptr_t
 * (0 = the legacy default stream).  "host" pointers are caller-owned CPU memory,
 * "dev" pointers are CUDA device memory on the handle's device.
 */
#ifndef DQNB200_H_
#define DQNB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DQNB_OK 0
#define DQNB_EINVAL (-1)   /* bad argument */
#define DQNB_ECUDA (-2)    /* CUDA runtime error (message has the cudaError string) */
#define DQNB_ENOMEM (-3)   /* host or device allocation failed */
#define DQNB_ESTATE (-4)   /* call not valid in the handle's current state */

typedef struct dqnb_replay dqnb_replay;
typedef struct dqnb_logits dqnb_logits;
typedef struct dqnb_learner dqnb_learner;

int dqnb_version(void);
const char* dqnb_last_error(void);
int dqnb_device_count(int* count);
/* Raw device buffers for callers without their own allocator (tests, non-Python hosts). */
int dqnb_dev_alloc(int device, size_t bytes, void** dev_ptr);
int dqnb_dev_free(int device, void* dev_ptr);
int dqnb_copy_h2d(void* dev_dst, const void* host_src, size_t bytes, uintptr_t stream);
int dqnb_copy_d2h(void* host_dst, const void* dev_src, size_t bytes, uintptr_t stream);
int dqnb_stream_sync(uintptr_t stream);

/* ------------------------------------------------------------------ replay frame store
 * Stands in for the item storage of TransitionReplay / ReservoirTransitionReplay
 * (replay.py:163-333) and CircularBuffer / MGSC*TransitionReplay storage
 * (replay_circular.py:90-146,568-664,1217-1312) for items shaped like
 * replay.Transition (replay.py:39-44): two HxWxS uint8 stacks + action, reward, discount.
 *
 * The store keeps single HxW frames ("planes") in an HBM pool, deduplicated across the
 * overlapping stacks of consecutive transitions, plus a per-slot table of the 2*S plane
 * ids (-1 = the all-zero plane the processor pads with, processors.py:57-69).  Which slot
 * an item goes to (FIFO ring position or reservoir index j) is decided by the host-side
 * mirror of the reference class, which also owns the NumPy random stream. */
int dqnb_replay_create(int64_t capacity, int height, int width, int stack, int64_t initial_planes,
                       int device, dqnb_replay** out);
int dqnb_replay_destroy(dqnb_replay* h);
/* replaces `add` storage (replay.py:184-193, :267-283): put one transition into `slot`,
 * releasing whatever the slot held.  s_tm1/s_t are host HxWxS uint8 (C order). */
int dqnb_replay_put(dqnb_replay* h, int64_t slot, const uint8_t* s_tm1, const uint8_t* s_t,
                    int32_t action, float reward, float discount, uintptr_t stream);
/* push staged puts to the device (also done implicitly by gather). */
int dqnb_replay_flush(dqnb_replay* h, uintptr_t stream);
/* replaces get()+np.stack of sample() (replay.py:195-206): gather+stack n slots into
 * device arrays s_tm1/s_t [n,H,W,S] u8, a [n] i32, r [n] f32, discount [n] f32. */
int dqnb_replay_gather(dqnb_replay* h, const int64_t* slots_host, int64_t n, uint8_t* s_tm1_dev,
                       uint8_t* s_t_dev, int32_t* a_dev, float* r_dev, float* disc_dev,
                       uintptr_t stream);
/* same, slot indices already on the device (int32), for use inside a captured CUDA graph. */
int dqnb_replay_gather_dev(dqnb_replay* h, const int32_t* slots_dev, int64_t n, uint8_t* s_tm1_dev,
                           uint8_t* s_t_dev, int32_t* a_dev, float* r_dev, float* disc_dev,
                           uintptr_t stream);
/* gather + device->host copy into caller (host) arrays; blocks until the data is there.
 * This is what the drop-in `sample()` returns as NumPy arrays. */
int dqnb_replay_gather_host(dqnb_replay* h, const int64_t* slots_host, int64_t n, uint8_t* s_tm1_host,
                            uint8_t* s_t_host, int32_t* a_host, float* r_host, float* disc_host,
                            uintptr_t stream);
/* tuning knob for the gather grid: CTAs per sampled transition (0 = automatic). */
int dqnb_replay_set_gather_splits(dqnb_replay* h, int splits);
/* counters: [0] planes in use, [1] pool capacity (planes), [2] puts, [3] planes uploaded,
 * [4] chained puts (s_tm1 matched the previous s_t), [5] pool growths, [6] bytes of HBM held */
int dqnb_replay_stats(dqnb_replay* h, int64_t out[8]);

/* Checkpoint of the whole store (get_state / set_state of the replay classes: replay.py:222-235,
 * replay_circular.py:132-146,643-660; the reference pickles one snappy-compressed item per slot): ONE blob -
 * header, host bookkeeping (plane ref-counts, free list, the chaining state), slot tables and the plane pool,
 * the device parts copied at PCIe rate.  Layout: int64 header fields [magic "DQNBRPLY", version, capacity, H, W,
 * S, plane_stride, pool_planes, n_free, have_last, byte offsets of {slot_ids int32[capacity][8], actions
 * int32[capacity], rewards f32, discounts f32, refcount int32[pool], free int32[n_free], last_ids int32[8],
 * last_st u8[H*W*S], planes u8[pool][plane_stride]}, total bytes, 4 counters].  import requires a store of the
 * same capacity and frame shape and replaces its contents.  Both synchronous. */
int dqnb_replay_export_size(dqnb_replay* h, int64_t* bytes);
int dqnb_replay_export(dqnb_replay* h, void* host, int64_t bytes, uintptr_t stream);
int dqnb_replay_import(dqnb_replay* h, const void* host, int64_t bytes, uintptr_t stream);

/* ------------------------------------------------------------------ learner
 * Stands in for the jitted `update` closure of Dqn / DoubleDqn (dqn/agent.py:85-119,
 * double_q/agent.py:85-121): Nature-CNN forward (networks.py:181-221,338-363) of the online
 * net on s_tm1 and the target net on s_t (+ the online net on s_t for double-Q), the rlax
 * q_learning / double_q_learning TD error with clip_gradient and l2_loss, reverse mode
 * through the online net, optax.rmsprop(centered=True) + apply_updates
 * (dqn/run_atari.py:208-213), and the target sync of dqn/agent.py:155-156.
 * Parameters live in one flat fp32 arena in haiku leaf order:
 * conv2_d{w[8,8,4,32],b} conv2_d_1{w[4,4,32,64],b} conv2_d_2{w[3,3,64,64],b}
 * linear{w[3136,512],b} linear_1{w[512,A], b[A] or b[1] when shared_bias}. */
#define DQNB_FP32 0   /* SIMT fp32 kernels; parity tolerance rel 1e-5 */
#define DQNB_TF32 2   /* tcgen05 tensor cores: kind::tf32 on the fp32 arrays (kind::f16 on fp16 copies of frames / post-ReLU
                         activations / filters for the forward convolutions - same 11-bit significand), fp32 accumulate
                         in TMEM; parity tolerance rel 1e-3.  (1 was a reserved bf16 value; never built.) */

typedef struct dqnb_learner_config {
  int32_t num_actions;
  int32_t shared_bias;      /* 1: single shared output bias (networks.py:120-134) */
  int32_t double_q;         /* 1: rlax.double_q_learning (double_q/agent.py:97-104) */
  int32_t batch_size;
  int32_t loss_mean_over;   /* rows the loss mean runs over; 0 = batch_size (DP ranks pass the global batch) */
  int32_t precision;        /* DQNB_FP32 or DQNB_TF32 */
  float learning_rate;      /* optax.rmsprop learning_rate */
  float decay;              /* optax.rmsprop decay */
  float eps;                /* optax.rmsprop eps (inside the rsqrt) */
  float grad_error_bound;   /* rlax.clip_gradient bound */
  int32_t defer_apply;      /* 1: update() stops after the gradient (DQNB_GRAD holds it); the optimiser
                               step runs in dqnb_learner_apply_grads - the seam for the data-parallel
                               all-reduce and for the MGSC weighted gradient */
  int32_t row_weighted;     /* 1: every row's clipped cotangent is multiplied by DQNB_ROW_WEIGHT[row] (initially
                               1) - with loss_mean_over = 1 the gradient is sum_i w_i grad(loss_i), the
                               vmapped weighted_grads of dqn_mgsc_batched/agent.py:171-176 as ONE batched pass.
                               2: importance-sampling weights of prioritized replay (prioritized/agent.py:85-115):
                               loss = mean(w_i * 0.5 td_i^2), so the weight multiplies the cotangent BEFORE
                               rlax.clip_gradient clips it: dq = -clip(w_i td_i / B) */
} dqnb_learner_config;

/* state arrays addressable by read/write */
#define DQNB_ONLINE 0
#define DQNB_TARGET 1
#define DQNB_MU 2
#define DQNB_NU 3
#define DQNB_GRAD 4       /* flat gradient of the last update */
#define DQNB_Q_TM1 5      /* [batch, A] online q(s_tm1) of the last update */
#define DQNB_Q_TARGET_T 6 /* [batch, A] target q(s_t) */
#define DQNB_Q_T 7        /* [batch, A] online q(s_t) (double-Q only) */
#define DQNB_TD 8         /* [batch] td errors */
#define DQNB_LOSS 9       /* [1] mean loss */
#define DQNB_ROW_WEIGHT 10 /* [batch] per-row cotangent weights (row_weighted learners; device-writable) */

int dqnb_learner_create(const dqnb_learner_config* cfg, int device, dqnb_learner** out);
int dqnb_learner_destroy(dqnb_learner* h);
int dqnb_learner_num_params(dqnb_learner* h, int64_t* n);
/* host <-> device copies of a state array (synchronous); n must equal the array length. */
int dqnb_learner_write(dqnb_learner* h, int which, const float* host, int64_t n, uintptr_t stream);
int dqnb_learner_read(dqnb_learner* h, int which, float* host, int64_t n, uintptr_t stream);
/* device pointer + length of a state array (library-owned; valid until destroy).  Lets the
 * caller hand the flat gradient arena to NCCL, or alias state into its own tensors. */
int dqnb_learner_device_ptr(dqnb_learner* h, int which, void** ptr, int64_t* n);
/* device -> device write of a state array (same rules as dqnb_learner_write; asynchronous). */
int dqnb_learner_write_dev(dqnb_learner* h, int which, const float* dev_src, int64_t n, uintptr_t stream);
/* optimizer.update + apply_updates (dqn/agent.py:115-116) on whatever DQNB_GRAD holds; the second
 * half of an update for learners created with defer_apply = 1. */
int dqnb_learner_apply_grads(dqnb_learner* h, uintptr_t stream);
/* ---- data-parallel learner (north_star: "large-batch data-parallel learner with an NCCL allreduce of the ~1.7M-
 * parameter gradient over NVLink"; the reference has no multi-device code, this is how its `update`,
 * dqn/agent.py:109-117, is spread over GPUs).  One process per GPU; every rank creates a learner with
 * batch_size = its share of the rows, loss_mean_over = the GLOBAL batch and defer_apply = 1, then:
 *   rank 0: dqnb_dp_unique_id(id) -> the 128 bytes reach the other ranks by any side channel
 *   all   : dqnb_dp_init(learner, id, rank, world)   (collective: blocks until every rank has called it)
 * From then on ONE dqnb_learner_update* call per rank is the whole data-parallel update: its CUDA graph runs
 * forward / TD / backward on the local rows, sums the split-K partials of [fc1.w, end) and all-reduces them
 * (ncclAllReduce, sum, fp32) on a side stream while the convolution backward still runs, all-reduces the
 * convolution gradients behind it, and applies the optimiser step.  Every rank must issue the same sequence of
 * updates.  NCCL is loaded with dlopen (the process's own libnccl.so.2 if one is mapped, e.g. PyTorch's; else
 * the system one; DQNB_NCCL_LIB overrides). */
int dqnb_dp_unique_id(uint8_t* id128);
int dqnb_dp_init(dqnb_learner* h, const uint8_t* id128, int rank, int world);
/* the two collectives of one update alone on `stream` (timing aid; contents of DQNB_GRAD are summed in place). */
int dqnb_dp_allreduce(dqnb_learner* h, uintptr_t stream);
int dqnb_dp_info(dqnb_learner* h, int* rank, int* world, int* nccl_version);

/* row_weighted learners, after an update: out_dev[i] = <v, w_i * d loss_i / d theta> for every row i of
 * that update's batch, v_dev a device vector laid out like the parameters.  Computed as the sum over
 * layers of <saved cotangent of row i, layer-local forward product of v> - four forward GEMMs and
 * one dot kernel instead of `batch` materialised per-sample gradients (the reference's
 * jax.vmap(weighted_grads), dqn_mgsc_batched/agent.py:171, holds M x 6.75 MB).  Asynchronous. */
int dqnb_learner_adjoint_dots(dqnb_learner* h, const float* v_dev, float* out_dev, uintptr_t stream);
/* Hessian-vector product out_dev = H v_dev, H the Jacobian of jax.grad(loss_fn) w.r.t. the online parameters at the
 * batch of this learner's LAST update (its td errors, clip decisions and ReLU masks are reused): what differentiating
 * through `jax.grad(loss_fn)(expected_online_params, ...)` needs in the second-order MGSC objective
 * (dqn_mgsc_batched_reservoir/agent.py:160-198, no stop_gradient on target_online_params).  Forward-over-reverse with
 * the layer GEMMs (rlax.clip_gradient's backward is differentiated: 1 inside the bounds, 0 outside).  Both vectors are
 * device memory laid out like the parameters.  Asynchronous. */
int dqnb_learner_hvp(dqnb_learner* h, const float* v_dev, float* out_dev, uintptr_t stream);
/* one `update` on a batch that is already in device memory (asynchronous). */
int dqnb_learner_update(dqnb_learner* h, const uint8_t* s_tm1_dev, const int32_t* a_dev,
                        const float* r_dev, const float* disc_dev, const uint8_t* s_t_dev,
                        uintptr_t stream);
/* one `update` on host arrays, the way the reference calls its jitted update with NumPy
 * transitions (dqn/agent.py:179-189): H2D copy through pinned staging, then the step. */
int dqnb_learner_update_host(dqnb_learner* h, const uint8_t* s_tm1_host, const int32_t* a_host,
                             const float* r_host, const float* disc_host, const uint8_t* s_t_host,
                             uintptr_t stream);
/* `_learn` (dqn/agent.py:179-189) with the sampled batch never leaving HBM: gather+stack the
 * given replay slots straight into the learner's batch buffers, then the step. */
int dqnb_learner_update_from_replay(dqnb_learner* h, dqnb_replay* replay, const int64_t* slots_host,
                                    uintptr_t stream);
/* the same with per-row weights for this batch (host float32 [batch]; learners created with row_weighted != 0):
 * the importance-sampling weights of `_learn` in prioritized/agent.py:215-236. */
int dqnb_learner_update_from_replay_weighted(dqnb_learner* h, dqnb_replay* replay, const int64_t* slots_host,
                                             const float* weights_host, uintptr_t stream);
/* target_params = online_params (dqn/agent.py:155-156). */
int dqnb_learner_sync_target(dqnb_learner* h, uintptr_t stream);
/* Q-values of n <= batch_size host observations under the online (0) or target (1) params;
 * the forward half of select_action (dqn/agent.py:121-129).  Synchronous. */
int dqnb_learner_q_values(dqnb_learner* h, int which_params, const uint8_t* s_host, int32_t n,
                          float* q_host, uintptr_t stream);
/* Diagnostics: per-kernel device times of one update (the graph's launches run eagerly on one
 * stream with a CUDA event after each), averaged over `iters` updates of the batch currently in
 * the learner's buffers; writes a text table "name  us" into `report`. */
int dqnb_learner_profile(dqnb_learner* h, int iters, char* report, size_t report_len, uintptr_t stream);
/* Diagnostics: begin/end (us, from %globaltimer) of every kernel of one update as the CUDA graph
 * really runs it; text table into `report`.  Advances the state by 3 updates. */
int dqnb_learner_timeline(dqnb_learner* h, char* report, size_t report_len, uintptr_t stream);
/* Diagnostics: `times` more updates on the batch that is already in the learner's buffers (graph replays
 * only: no copies, no gather) - the period of back-to-back updates minus the timeline's span is the
 * graph-to-graph gap.  Advances the learner state like `times` updates.  Asynchronous. */
int dqnb_learner_relaunch(dqnb_learner* h, int times, uintptr_t stream);
/* Loss read-back that neither drains the stream nor uses a copy engine: every update writes
 * (sequence number, mean loss) into a pinned host ring of 8 entries.  `post` notes the sequence
 * number of the update launched last; `take` polls host memory until the OLDEST posted entry has
 * landed and returns it (per-step statistics, consumed one step behind, learner never idle).
 * At most 8 posts may be outstanding, and a post must be taken before 8 further updates are
 * launched (overrun is reported as an error); `take` with none outstanding is an error. */
int dqnb_learner_loss_post(dqnb_learner* h, uintptr_t stream);
int dqnb_learner_loss_take(dqnb_learner* h, float* loss);
/* number of kernel launches (graph nodes) one update issues. */
int dqnb_learner_launches_per_update(dqnb_learner* h, int64_t* n);

/* the configuration and device a learner was created with. */
int dqnb_learner_get_config(dqnb_learner* h, dqnb_learner_config* cfg, int* device);

/* ---------------------------------------------------------------------------------------------
 * MGSC meta_update (dqn_mgsc_batched/agent.py:160-220; meta optimiser optax.adam,
 * dqn_mgsc_batched/run_atari.py:241-243, state shape (meta_batch_size,) agent.py:80).
 * Reads the learner's online/target params and RMSProp state without advancing them (virtual
 * steps), returns the updated logits of the meta batch and advances its own Adam state.
 * The learner must use the q_learning loss (double_q = 0).  Synchronous.
 * ------------------------------------------------------------------------------------------- */
typedef struct dqnb_mgsc dqnb_mgsc;
typedef struct dqnb_mgsc_config {
  int32_t meta_batch_size;   /* M */
  float learning_rate;       /* optax.adam learning_rate */
  float b1, b2, eps;         /* optax.adam defaults 0.9, 0.999, 1e-8 */
  int32_t second_order;      /* 0: dqn_mgsc_batched (stop_gradient on target_online_params, agent.py:189-191);
                                1: dqn_mgsc_batched_reservoir/agent.py:160-198 (no stop_gradient: the meta loss is
                                differentiated through the gradient at the expected parameters - dqnb_learner_hvp) */
} dqnb_mgsc_config;
#define DQNB_MGSC_ADAM_M 0      /* [M] Adam first moment (read/write) */
#define DQNB_MGSC_ADAM_V 1      /* [M] Adam second moment (read/write) */
#define DQNB_MGSC_P 2           /* [M] softmax(logits) of the last call */
#define DQNB_MGSC_DL_DP 3       /* [M] dL/dp_i of the last call */
#define DQNB_MGSC_G 4           /* [params] sum_i p_i g_i */
#define DQNB_MGSC_V 5           /* [params] dL/dG */
#define DQNB_MGSC_THETA_EXP 6   /* [params] expected online params */
int dqnb_mgsc_create(dqnb_learner* learner, const dqnb_mgsc_config* cfg, dqnb_mgsc** out);
int dqnb_mgsc_destroy(dqnb_mgsc* h);
/* meta batch = M replay slots gathered on the device (replay_circular.py:1270-1275), online
 * transition and logits from host memory; new_logits[M] and *meta_loss (nullable) to host. */
int dqnb_mgsc_meta_update(dqnb_mgsc* h, dqnb_replay* replay, const int64_t* slots_host, const float* logits,
                          const uint8_t* on_s_tm1, int32_t on_a, float on_r, float on_disc,
                          const uint8_t* on_s_t, float* new_logits, float* meta_loss, uintptr_t stream);
/* same with the meta batch in host arrays ([M,84,84,4] uint8 stacks, [M] scalars). */
int dqnb_mgsc_meta_update_host(dqnb_mgsc* h, const uint8_t* s_tm1, const int32_t* a, const float* r,
                               const float* disc, const uint8_t* s_t, const float* logits,
                               const uint8_t* on_s_tm1, int32_t on_a, float on_r, float on_disc,
                               const uint8_t* on_s_t, float* new_logits, float* meta_loss, uintptr_t stream);
int dqnb_mgsc_read(dqnb_mgsc* h, int which, float* host, int64_t n);
int dqnb_mgsc_write(dqnb_mgsc* h, int which, const float* host, int64_t n);
/* Test hook: dqnb_learner_read on the internal batch-of-M (which_learner 0) or batch-of-1 (1) learner, as the
 * last meta_update left it (activations / td errors the parity tests condition the oracle on). */
int dqnb_mgsc_read_learner(dqnb_mgsc* h, int which_learner, int which, float* host, int64_t n);
/* Adam step count: *get receives it (nullable), *set overwrites it (nullable). */
int dqnb_mgsc_adam_count(dqnb_mgsc* h, int32_t* get, const int32_t* set);

/* ---------------------------------------------------------------------------------------------
 * MGSC logit ring in device memory (replay_circular.py:148-248 CircularLogitBuffer; the logits of
 * MGSCReservoirDistribution :500-560).  The array of `capacity` float32 logits (-inf = empty) lives
 * in HBM; ring heads, sizes and the NumPy Generator stay with the caller.  Indices are ABSOLUTE.
 * ------------------------------------------------------------------------------------------- */
typedef struct dqnb_logits dqnb_logits;
int dqnb_logits_create(int64_t capacity, int device, dqnb_logits** out);
int dqnb_logits_destroy(dqnb_logits* h);
/* logits[index] = size == 0 ? 0 : logsumexp(all logits) - log(size)  (add(), :166-179; reservoir
 * replace :526-533 after the slot was set to -inf).  Asynchronous, no host round trip. */
int dqnb_logits_fill(dqnb_logits* h, int64_t index, int64_t size, uintptr_t stream);
int dqnb_logits_set(dqnb_logits* h, const int64_t* indices, const float* values, int64_t n, uintptr_t stream);
int dqnb_logits_get(dqnb_logits* h, const int64_t* indices, float* values, int64_t n, uintptr_t stream);
/* Generator.choice(capacity, n, p=softmax(logits)) for the n doubles the caller drew with
 * Generator.random(n) (what choice() consumes, :205-217): float32 softmax, float64 CDF,
 * searchsorted(side='right').  Synchronous. */
int dqnb_logits_sample(dqnb_logits* h, const double* uniforms, int64_t n, int64_t* indices, uintptr_t stream);
int dqnb_logits_lse(dqnb_logits* h, float* lse, uintptr_t stream);
/* whole-array copies for get_state / set_state / `_logits` pokes; n must equal capacity. */
int dqnb_logits_read(dqnb_logits* h, float* host, int64_t n);
int dqnb_logits_write(dqnb_logits* h, const float* host, int64_t n);
/* Allocation-free logsumexp.  No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  *lse_dev = logsumexp(logits_dev[0..n)) in float32, computed by the same kernels as dqnb_logits_lse, with
 * their scratch (per-block partials) in a fixed file-scope device array: the call allocates and frees nothing.
 * Limit: n <= 4 194 304 (rejected with DQNB_EINVAL and a message beyond it); the 16.6 KB array exists on every GPU
 * that loads the library until the process exits.  All calls share the array: calls that overlap on the GPU
 * (different streams without ordering between them) race - serialise them.  Both pointers are device memory on
 * `device`; asynchronous on `stream`. */
int dqnb_logsumexp_noalloc(int device, const float* logits_dev, int64_t n, float* lse_dev, uintptr_t stream);

/* ---------------------------------------------------------------------------------------------
 * Prioritized replay sum tree in device memory (replay.py:379-559 SumTree; used by PrioritizedDistribution
 * :562-784).  storage[2 * first_leaf] float64, node i = storage[2i] + storage[2i+1], root at 1, leaves from
 * first_leaf (a power of two, the reference's `capacity`): the reference's layout and the reference's float64
 * arithmetic, so nodes and query results are bit-identical.  Which leaf an id owns, sizes and the NumPy random
 * stream stay with the caller.
 * ------------------------------------------------------------------------------------------- */
typedef struct dqnb_sumtree dqnb_sumtree;
int dqnb_sumtree_create(int64_t first_leaf, int device, dqnb_sumtree** out);
int dqnb_sumtree_destroy(dqnb_sumtree* h);
/* SumTree.set (replay.py:402-414): leaves[indices] = values (duplicates: last wins), ancestors re-summed.
 * Values must be finite and >= 0.  Asynchronous. */
int dqnb_sumtree_set(dqnb_sumtree* h, const int64_t* indices, const double* values, int64_t n, uintptr_t stream);
/* SumTree.query (replay.py:424-440): for each target the smallest leaf with target < inclusive cumulative sum.
 * scale_by_root = 1: the inputs are uniforms in [0,1) and target = uniform * root is formed on the device
 * (PrioritizedDistribution.sample, replay.py:697-699).  *root receives storage[1] (nullable).  Fails with the
 * reference's message when a target is outside [0, root).  Synchronous. */
int dqnb_sumtree_query(dqnb_sumtree* h, const double* targets_or_uniforms, int64_t n, int scale_by_root,
                       int64_t* indices, double* root, uintptr_t stream);
/* whole-storage copies (n = 2 * first_leaf): get_state / set_state / resize / set_all. */
int dqnb_sumtree_read(dqnb_sumtree* h, double* storage_host, int64_t n);
int dqnb_sumtree_write(dqnb_sumtree* h, const double* storage_host, int64_t n);

/* Test hook (not part of the drop-in surface): C[M,N] = A*B^T through the TF32 slab kernel on fp32 device matrices
 * (tcgen05.mma.kind::tf32); *_mn_major selects [K,M]/[K,N] storage; M, N, K multiples of 4. */
int dqnb_selftest_tf32_gemm(const float* a_dev, const float* b_dev, float* c_dev, int M, int N, int K,
                            int a_mn_major, int b_mn_major, int bn, int splits, uintptr_t stream);

#ifdef __cplusplus
}
#endif
#endif /* DQNB200_H_ */
