#!/usr/bin/env python
"""Headline benchmark: learner updates/s (double-Q Nature-DQN, batch 32, 1M-transition HBM
replay, synthetic 84x84x4 uint8 frames) + replay gather GB/s.  See DESIGN.md "Measurement".

  python bench.py --gpus N --steps K --warmup W            # this repo (CUDA path)
  python bench.py --impl reference --gpus N --steps K ...  # CPU port of the reference path

One JSON line on stdout (rank 0).  N>1: launched by torchrun, one process per GPU, each rank
an independent agent replica (own replay, params, RNG) - the way the reference scales (SLURM
arrays); no data-path collective; value = total updates of all ranks / max-over-ranks time.
"""
import argparse
import collections
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

NUM_ACTIONS = 6
BATCH = 32
ADDS_PER_LEARN = 4                 # learn_period 16 frames / action repeat 4 (dqn/agent.py:152)
TARGET_SYNC_UPDATES = 7500         # 1.2e5 frames / 16 (double_q/run_atari.py:86)
PLANE = 84 * 84
GATHER_BYTES_PER_TRANSITION = 13 * PLANE + 12     # 5 planes read + 8 written + scalars
UPDATE_BYTES = 49.05e6             # params/opt-state touched once + batch
NCU_GATHER_TRAFFIC_BYTES = 308.5e6 # dram__bytes_read.sum + dram__bytes_write.sum of gather_stack4 at n=4096 (r01c)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--capacity", type=int, default=1_000_000)
    ap.add_argument("--gather-n", type=int, default=4096)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--mode", default="replicas", choices=["replicas", "dp"],
                    help="N>1: independent agent replicas (weak scaling, the reference's way) or one "
                         "data-parallel learner on a fixed global batch (strong scaling, NCCL all-reduce)")
    ap.add_argument("--global-batch", type=int, default=4096, help="dp mode: global batch")
    ap.add_argument("--precision", default="tf32", choices=["tf32", "fp32"],
                    help="arithmetic of the conv/dense GEMMs: tf32 = tcgen05 tensor cores, fp32 = SIMT")
    return ap.parse_args()


class EpisodeSource:
    """Fast synthetic transitions with the reference stack layout (A000, AB00, ... ) and
    chained observations (s_tm1 of a transition *is* the previous s_t object)."""

    def __init__(self, seed, block=2048):
        self.rng = np.random.default_rng(seed)
        self.block = block
        self.frames = None
        self.pos = block
        self.left = 0
        self.prev = None
        self.window = None

    def _frame(self):
        if self.pos == self.block:
            self.frames = self.rng.integers(0, 256, size=(self.block, 84, 84), dtype=np.uint8)
            self.pos = 0
        f = self.frames[self.pos]
        self.pos += 1
        return f

    def _obs(self, frame):
        self.window = (self.window + [frame])[-4:]
        obs = np.zeros((84, 84, 4), dtype=np.uint8)
        for c, f in enumerate(self.window):
            obs[:, :, c] = f
        return obs

    def next(self):
        """-> (s_tm1, a, r, discount, s_t)"""
        if self.left == 0:
            self.left = int(self.rng.integers(50, 2000))
            self.window = []
            self.prev = self._obs(self._frame())
        self.left -= 1
        s_t = self._obs(self._frame())
        u = self.rng.random(2)
        r = -1.0 if u[0] < 0.05 else (1.0 if u[0] > 0.95 else 0.0)
        d = 0.0 if (self.left == 0 or u[1] < 0.01) else 0.99
        tr = (self.prev, int(self.rng.integers(0, NUM_ACTIONS)), r, d, s_t)
        self.prev = s_t
        return tr


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device, self.proc, self.lines = device, None, []

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.device), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=lambda: self.lines.extend(self.proc.stdout), daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def __exit__(self, *exc):
        if self.proc:
            time.sleep(0.12)
            self.proc.terminate()
            self.thread.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1]))
            except ValueError:
                continue
            for n, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm)}


def cpu_reference_arm(seconds, steps=None, warmup=3, capacity=4096):
    """The reference path on host cores: replay.sample(32) (oracle restatement of replay.py)
    + `update` (torch-CPU fp32 port, all threads).  Bounded sample of the same workload."""
    from oracle import replay_oracle as ro
    from oracle import torch_port
    port = torch_port.make(NUM_ACTIONS, shared_bias=True, double_q=True, threads=os.cpu_count())
    rep = ro.FifoReplay(capacity, ro.Transition(None, None, None, None, None), np.random.RandomState(0))
    src = EpisodeSource(1, block=256)
    for _ in range(capacity):
        rep.add(ro.Transition(*src.next()))

    def one_step():
        for _ in range(ADDS_PER_LEARN):
            rep.add(ro.Transition(*src.next()))
        s0, a, r, d, s1 = rep.sample(BATCH)
        return port.update(s0, a, r, d, s1)

    for _ in range(warmup):
        one_step()
    t0 = time.perf_counter()
    n = 0
    while True:
        one_step()
        n += 1
        el = time.perf_counter() - t0
        if (steps is not None and n >= steps) or (steps is None and el >= seconds) or el > 240:
            break
    return {"value": n / el, "unit": "updates/s", "cores": port.threads, "kind": "port",
            "sample": "%d updates of batch %d (4 adds + sample(32) from a %d-transition host FIFO replay + "
                      "torch-CPU fp32 double-Q update), %.1f s" % (n, BATCH, capacity, el),
            "steps": n, "seconds": el}


def run_reference(args, rank, world):
    if rank != 0:
        return
    res = cpu_reference_arm(args.cpu_seconds, steps=args.steps, warmup=max(3, min(args.warmup, 10)))
    line = {
        "impl": "reference", "metric": "learner_updates_per_s", "value": res["value"], "unit": "updates/s",
        "n_gpus": args.gpus, "steps": res["steps"], "warmup": max(3, min(args.warmup, 10)),
        "ms_per_step": 1e3 / res["value"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": workload_config(args, res_capacity=4096),
        "cpu_baseline": {k: res[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": res["value"], "unit": "updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def workload_config(args, res_capacity=None):
    return {"workload": "double_q learner, Nature CNN, batch 32, %d-transition FIFO replay, synthetic 84x84x4 "
                        "uint8 frames, %d actions" % (res_capacity or args.capacity, NUM_ACTIONS),
            "batch": BATCH, "replay_capacity": res_capacity or args.capacity, "adds_per_learn": ADDS_PER_LEARN,
            "target_sync_every_updates": TARGET_SYNC_UPDATES, "parallelism": "replicas x%d" % args.gpus,
            "l2": "replay pool (7 GB) >> L2, random slots every step; params/opt state L2-resident as in training"}


def run_ours(args, rank, world, local_rank):
    import torch
    from dqn_mgsc_zoo_b200 import learner as lrn
    from dqn_mgsc_zoo_b200 import replay as rp
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- build the workload: full replay + learner with 100 warm updates worth of state
    t_fill = time.perf_counter()
    replay = rp.TransitionReplay(args.capacity, rp.Transition(None, None, None, None, None),
                                 np.random.RandomState(rank), device=local_rank)
    src = EpisodeSource(1000 + rank)
    for _ in range(args.capacity):
        replay.add(rp.Transition(*src.next()))
    store = replay.frame_store
    store.flush()
    t_fill = time.perf_counter() - t_fill
    L = lrn.Learner(NUM_ACTIONS, BATCH, shared_bias=True, double_q=True, precision=args.precision,
                    device=local_rank)
    from dqn_mgsc_zoo_b200 import networks
    init = networks.init_params(np.random.default_rng(rank), NUM_ACTIONS, shared_bias=True)
    L.set_params(init, lrn.ONLINE)
    L.set_params(init, lrn.TARGET)

    state = {"updates": 0}

    def device_step():
        slots = replay.sample_slots(BATCH)
        L.update_from_replay(store, slots)
        state["updates"] += 1
        if state["updates"] % TARGET_SYNC_UPDATES == 0:
            L.sync_target()

    def e2e_step():
        for _ in range(ADDS_PER_LEARN):
            replay.add(rp.Transition(*pending.popleft()))   # in arrival order: s_tm1 is the previous s_t
        device_step()
        return L.loss_lagged()   # queue this step's loss D2H, return the previous step's (already landed)

    def e2e_step_sync():
        for _ in range(ADDS_PER_LEARN):
            replay.add(rp.Transition(*pending.popleft()))   # in arrival order: s_tm1 is the previous s_t
        device_step()
        return L.loss()          # blocking D2H read of this step's loss

    def timed(fn, steps):
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        ev0.record()
        for _ in range(steps):
            fn()
        ev1.record()
        barrier()
        ms = ev0.elapsed_time(ev1)
        if dist is not None:
            t = torch.tensor([ms], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    for _ in range(max(3, args.warmup)):
        device_step()
    with ClockSampler(local_rank) as clocks:
        ms = timed(device_step, args.steps)
    value = world * args.steps / (ms * 1e-3)

    # ---- e2e: new transitions arrive from host memory every step, loss is read back
    e2e_steps = min(args.steps, 1000)
    sync_steps = min(args.steps, 300)
    pending = collections.deque(src.next() for _ in range(ADDS_PER_LEARN * (e2e_steps + sync_steps + 10)))
    for _ in range(5):
        e2e_step()
    up0 = store.stats()["planes_uploaded"]
    ms_e2e = timed(e2e_step, e2e_steps)      # the closing barrier waits for the last step's update
    last_loss = L.loss_take()                 # drain: the final step's loss
    planes = store.stats()["planes_uploaded"] - up0
    h2d = planes * PLANE / e2e_steps + ADDS_PER_LEARN * 64 + BATCH * 8
    for _ in range(3):
        e2e_step_sync()
    ms_sync = timed(e2e_step_sync, sync_steps)
    e2e = {"value": world * e2e_steps / (ms_e2e * 1e-3), "unit": "updates/s",
           "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": 4, "steps": e2e_steps,
           "what": "per step: 4 replay.add(host transition) [H2D of the new frames via pinned staging] + "
                   "sample_slots + update_from_replay + D2H of the step's loss; the loss copy is queued "
                   "behind the update and consumed one step later, so host work overlaps the device",
           "blocking_loss_updates_per_s": world * sync_steps / (ms_sync * 1e-3),
           "last_loss": last_loss}

    # ---- literal seam: update(host NumPy batch), 3.6 MB H2D per step like the reference's jit dispatch
    hb = store.gather_host(replay.sample_slots(BATCH))
    for _ in range(3):
        L.update_host(*hb)
    hsteps = min(args.steps, 300)
    ms_host = timed(lambda: (L.update_host(*hb), L.loss()), hsteps)
    e2e["host_batch_updates_per_s"] = world * hsteps / (ms_host * 1e-3)
    e2e["host_batch_h2d_bytes_per_step"] = 2 * BATCH * 28224 + 12 * BATCH

    # ---- replay gather throughput at n transitions per launch (HBM roofline kernel)
    n = args.gather_n
    bufs = [torch.empty((n, 84, 84, 4), dtype=torch.uint8, device="cuda") for _ in range(2)]
    ai = torch.empty(n, dtype=torch.int32, device="cuda")
    rf, df = torch.empty(n, device="cuda"), torch.empty(n, device="cuda")
    rs = np.random.RandomState(99 + rank)
    # slot lists pre-uploaded (int32, device): the timed region is the gather kernel alone
    slot_sets = [torch.from_numpy(rs.randint(args.capacity, size=n).astype(np.int32)).cuda() for _ in range(24)]

    def gather_step(i=[0]):
        store.gather_dev(slot_sets[i[0] % len(slot_sets)].data_ptr(), n, bufs[0].data_ptr(), bufs[1].data_ptr(),
                         ai.data_ptr(), rf.data_ptr(), df.data_ptr())
        i[0] += 1
    for _ in range(4):
        gather_step()
    greps = 20
    ms_g = timed(gather_step, greps) / greps
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak, peak_src = (peaks["hbm_gbs"], "measured") if "hbm_gbs" in peaks else (6650.0, "fallback")
    achieved = n * GATHER_BYTES_PER_TRANSITION / (ms_g * 1e-3) / 1e9
    roofline = {"kernel": "gather_stack4", "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": NCU_GATHER_TRAFFIC_BYTES, "peak_source": peak_src,
                "transitions_per_launch": n, "us_per_launch": ms_g * 1e3,
                "algorithmic_bytes_per_launch": n * GATHER_BYTES_PER_TRANSITION,
                "note": "CUDA events around %d back-to-back launches on the launching stream, slot lists already on "
                        "the device; 24 different random slot lists over a %.1f GB plane pool (>> L2). traffic = "
                        "dram read+write per launch from profiles/r01c_ncu_gather_full.csv (ncu --set full, "
                        "60k-transition pool; write-back of the last lines is still in L2 when the kernel ends)"
                        % (greps, store.stats()["hbm_bytes"] / 1e9)}
    ms_upd = ms / args.steps
    roofline_update = {"bound": "hbm", "achieved": UPDATE_BYTES / (ms_upd * 1e-3) / 1e9, "peak": peak,
                       "unit": "GB/s", "frac": UPDATE_BYTES / (ms_upd * 1e-3) / 1e9 / peak,
                       "note": "whole update graph vs 49.05 MB touched-once bytes; batch 32 is latency bound"}

    # ---- the same step with the GEMMs in plain fp32 (SIMT path), for the record
    other = {}
    if args.precision == "tf32":
        L32 = lrn.Learner(NUM_ACTIONS, BATCH, shared_bias=True, double_q=True, precision="fp32", device=local_rank)
        L32.set_params(init, lrn.ONLINE)
        L32.set_params(init, lrn.TARGET)

        def fp32_step():
            L32.update_from_replay(store, replay.sample_slots(BATCH))
        for _ in range(10):
            fp32_step()
        n32 = min(args.steps, 300)
        other["fp32_simt_updates_per_s"] = world * n32 / (timed(fp32_step, n32) * 1e-3)

    if rank != 0:
        return
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        res = cpu_reference_arm(args.cpu_seconds)
        cpu = {k: res[k] for k in ("value", "unit", "cores", "kind", "sample")}
    launches = L.launches_per_update() + 1
    line = {
        "metric": "learner_updates_per_s", "value": value, "unit": "updates/s", "n_gpus": world,
        "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32" if args.precision == "fp32" else args.precision, "data": "synthetic",
        "config": workload_config(args), "clocks": clocks.summary(), "e2e": e2e,
        "gpu_launches": int(launches * args.steps), "launches_per_update": int(launches),
        "roofline": roofline, "roofline_update": roofline_update, "cpu_baseline": cpu,
        "other_precisions": other, "replay_gather_gbs": achieved, "replay_fill_seconds": t_fill, "replay_stats": store.stats(),
    }
    emit(line)
    if dist is not None:
        dist.destroy_process_group()


UPDATE_FLOP_PER_SAMPLE = 86_906_880     # double-Q update, A=6


def run_dp(args, rank, world, local_rank):
    """one learner, fixed global batch, rows sharded over the ranks, one NCCL
    all-reduce of the flat 6.75 MB gradient arena per update.  Replay replicated on every rank."""
    import torch
    import torch.distributed as dist
    from dqn_mgsc_zoo_b200 import learner as lrn
    from dqn_mgsc_zoo_b200 import networks, parallel
    from dqn_mgsc_zoo_b200 import replay as rp
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    G = args.global_batch
    capacity = min(args.capacity, 200_000)     # replicated ring; sampling pattern, not size, matters here
    replay = rp.TransitionReplay(capacity, rp.Transition(None, None, None, None, None),
                                 np.random.RandomState(0), device=local_rank)      # same seed: replicas
    src = EpisodeSource(1000)
    for _ in range(capacity):
        replay.add(rp.Transition(*src.next()))
    store = replay.frame_store
    store.flush()
    dp = parallel.make_nccl_learner(NUM_ACTIONS, G, local_rank, rank, world, shared_bias=True, double_q=True,
                                    precision=args.precision)
    L = dp.learner
    init = networks.init_params(np.random.default_rng(0), NUM_ACTIONS, shared_bias=True)
    L.set_params(init, lrn.ONLINE)
    L.set_params(init, lrn.TARGET)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def device_step():
        dp.update_from_replay(store, replay.sample_slots(G))      # identical draw on every rank

    pending = collections.deque()

    def e2e_step():
        for _ in range(ADDS_PER_LEARN):
            replay.add(rp.Transition(*pending.popleft()))   # in arrival order: s_tm1 is the previous s_t
        device_step()
        return L.loss_lagged()

    def timed(fn, steps):
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        ev0.record()
        for _ in range(steps):
            fn()
        ev1.record()
        barrier()
        ms = ev0.elapsed_time(ev1)
        if world > 1:
            t = torch.tensor([ms], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    steps = min(args.steps, 300)
    for _ in range(max(3, min(args.warmup, 10))):
        device_step()
    with ClockSampler(local_rank) as clocks:
        ms = timed(device_step, steps)
    value = steps / (ms * 1e-3)
    e2e_steps = min(steps, 100)
    pending.extend(src.next() for _ in range(ADDS_PER_LEARN * (e2e_steps + 5)))
    for _ in range(3):
        e2e_step()
    ms_e2e = timed(e2e_step, e2e_steps)
    L.loss_take()
    params = L.read(lrn.ONLINE)
    digest = float(np.abs(params).sum())
    same = True
    if world > 1:     # replicas must hold bit-identical parameters after the run
        t = torch.tensor([digest], device="cuda", dtype=torch.float64)
        lo, hi = t.clone(), t.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        same = bool(lo.item() == hi.item())
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        tf = UPDATE_FLOP_PER_SAMPLE * G * value / 1e12
        peak = peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops", 1398.4))
        line = {
            "metric": "learner_updates_per_s", "value": value, "unit": "updates/s", "n_gpus": world, "steps": steps,
            "warmup": max(3, min(args.warmup, 10)), "ms_per_step": ms / steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32" if args.precision == "fp32" else args.precision,
            "data": "synthetic",
            "config": {"workload": "data-parallel double_q learner, Nature CNN, GLOBAL batch %d (%d per GPU), "
                                   "replicated %d-transition FIFO replay, one NCCL all-reduce of the 6.75 MB "
                                   "gradient arena per update" % (G, G // world, capacity),
                       "global_batch": G, "parallelism": "dp%d" % world,
                       "l2": "random slots from a %.1f GB plane pool every step" % (capacity * 7056 / 1e9)},
            "clocks": clocks.summary(), "samples_per_s": value * G,
            "e2e": {"value": e2e_steps / (ms_e2e * 1e-3), "unit": "updates/s",
                    "h2d_bytes_per_step": ADDS_PER_LEARN * (PLANE + 64) + G * 8, "d2h_bytes_per_step": 4,
                    "what": "4 replicated replay.add + sample_slots(global batch) + sharded update + all-reduce "
                            "+ apply + lagged loss read"},
            "gpu_launches": int((L.launches_per_update() + 2) * steps),
            "roofline": {"bound": "tensor", "achieved": tf / world, "peak": peak, "unit": "TFLOP/s",
                         "frac": tf / world / peak, "traffic": None,
                         "note": "per-GPU algorithmic flops (86.9 MFLOP x rows) / step time vs measured dense "
                                 "bf16 peak; kind::tf32 nominal peak is half of bf16"},
            "cpu_baseline": None, "replicas_identical": same,
        }
        emit(line)
    if world > 1:
        dist.destroy_process_group()


_JSON_OUT = None


def emit(line):
    """The one JSON line, on the process's ORIGINAL stdout."""
    out = _JSON_OUT if _JSON_OUT is not None else sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    global _JSON_OUT
    args = parse_args()
    # Libraries print to fd 1 behind Python's back (NCCL's version banner came out ahead of the JSON line
    # whatever NCCL_DEBUG_FILE said): keep the original stdout for the JSON line only and point fd 1 at
    # stderr for everything else.
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    elif args.mode == "dp":
        run_dp(args, rank, world, local_rank)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
