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


def ncu_gather_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per gather_stack4 launch (n = 4096) from the newest
    `profiles/*ncu_gather*.csv` (an `ncu --set full` capture of scripts/ncu_target.py).  -> (bytes, file)"""
    import csv
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "*ncu_gather*.csv")))
    for path in reversed(files):
        try:
            rows = [r for r in csv.DictReader(open(path)) if "gather_stack4" in r.get("kernel", "")
                    and float(r.get("launch__grid_size []", 0) or 0) >= 1024]      # the n = 4096 launches only
            rd = [float(r["dram__bytes_read.sum [Mbyte]"]) for r in rows]
            wr = [float(r["dram__bytes_write.sum [Mbyte]"]) for r in rows]
            if rd:
                return (float(np.median(rd)) + float(np.median(wr))) * 1e6, os.path.basename(path)
        except (OSError, KeyError, ValueError):
            continue
    return None, None


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--capacity", type=int, default=1_000_000)
    ap.add_argument("--gather-n", type=int, default=4096)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--cpu-capacity", type=int, default=None,
                    help="CPU arm: replay capacity (default: 1M when the host RAM holds it)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--windows", type=int, default=25,
                    help="timed windows of --steps updates each; the median window is reported")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the reservoir / MGSC / data-parallel sub-records (C3, C4, C5-ii)")
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


def _host_ram_free_bytes():
    try:
        import psutil
        return int(psutil.virtual_memory().available)
    except Exception:
        return 0


def cpu_reference_arm(seconds, steps=None, warmup=3, capacity=None, windows=1):
    """The reference path on the host cores, same workload as the CUDA arm:
      replay half  = the reference's OWN TransitionReplay.add / sample (dqn_zoo/replay.py:163-243), loaded
                     unmodified from /root/reference or its staged copy oracle/_ref (oracle/build_ref.py) with
                     encoder=decoder=None (no snappy here - this favours the baseline); the oracle restatement
                     only if neither exists ("port");
      learner half = torch-CPU fp32 port of `update` (oracle/torch_port.py), all host threads.
    The replay is as large as the host holds (1 M transitions = 28 GB of uint8 stacks when RAM allows).  The arm
    is warmed until two consecutive 20-update windows agree within 3 %, then `windows` windows of `steps` updates
    are timed and the median window is reported (or, with steps=None, everything inside `seconds`)."""
    from oracle import ref_shim, torch_port
    port = torch_port.make(NUM_ACTIONS, shared_bias=True, double_q=True, threads=os.cpu_count())
    want = capacity or 1_000_000
    per_item = 28224 + 400                    # one new (84,84,4) stack per transition (s_tm1 is the previous s_t object)
    free = _host_ram_free_bytes()
    if capacity is None and free and want * per_item > 0.5 * free:
        want = max(4096, int(0.5 * free / per_item))
    if ref_shim.available():
        ref = ref_shim.load("replay")
        rep = ref.TransitionReplay(want, ref.Transition(None, None, None, None, None), np.random.RandomState(0))
        make, kind = ref.Transition, "ref+port"
    else:
        from oracle import replay_oracle as ro
        rep = ro.FifoReplay(want, ro.Transition(None, None, None, None, None), np.random.RandomState(0))
        make, kind = ro.Transition, "port"
    src = EpisodeSource(1, block=256)
    t_fill = time.perf_counter()
    for _ in range(want):
        rep.add(make(*src.next()))
    t_fill = time.perf_counter() - t_fill

    def one_step():
        for _ in range(ADDS_PER_LEARN):
            rep.add(make(*src.next()))
        s0, a, r, d, s1 = rep.sample(BATCH)
        return port.update(s0, a, r, d, s1)

    def window(n):
        t0 = time.perf_counter()
        for _ in range(n):
            one_step()
        return time.perf_counter() - t0

    for _ in range(warmup):
        one_step()
    prev, warm_windows = window(20), 1
    while warm_windows < 40:
        cur = window(20)
        warm_windows += 1
        if abs(cur - prev) <= 0.03 * min(cur, prev):
            break
        prev = cur
    if steps is not None:
        times = [window(steps) for _ in range(max(1, windows))]
        el, n = float(np.median(times)), steps
    else:
        t0, n = time.perf_counter(), 0
        while True:
            one_step()
            n += 1
            el = time.perf_counter() - t0
            if el >= seconds or el > 240:
                break
    return {"value": n / el, "unit": "updates/s", "cores": port.threads, "kind": kind,
            "sample": "%d updates of batch %d (4 adds + sample(32) from a %d-transition host FIFO replay [%s, %.0f s "
                      "fill] + torch-CPU fp32 double-Q update), %.1f s%s, after %d warm-up windows of 20"
                      % (n, BATCH, want, "reference replay.py, unmodified" if kind == "ref+port" else "oracle port",
                         t_fill, el, " (median of %d windows)" % windows if steps is not None and windows > 1 else "",
                         warm_windows),
            "steps": n, "seconds": el, "replay_capacity": want}


REF_WINDOWS = 15


def run_reference(args, rank, world):
    if rank != 0:
        return
    res = cpu_reference_arm(args.cpu_seconds, steps=args.steps, warmup=max(3, min(args.warmup, 10)),
                            capacity=args.cpu_capacity, windows=REF_WINDOWS)
    cfg = workload_config(args, res_capacity=res["replay_capacity"])
    # one CPU process on rank 0 whatever N is: do not read value(N GPUs) / this as a same-N ratio
    cfg["parallelism"] = "cpu x1 (one process, %d threads%s)" % (res["cores"], ", rank 0 of %d" % world if world > 1 else "")
    cfg["l2"] = "host replay of %.1f GB >> CPU caches, random ids every step" % (res["replay_capacity"] * 28224 / 1e9)
    line = {
        "impl": "reference", "metric": "learner_updates_per_s", "value": res["value"], "unit": "updates/s",
        "n_gpus": args.gpus, "steps": res["steps"], "warmup": max(3, min(args.warmup, 10)),
        "ms_per_step": 1e3 / res["value"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": cfg,
        "timing": "median of %d consecutive windows of `steps` updates" % REF_WINDOWS,
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


def pin_to_cores(local_rank, world):
    """One slice of the host cores per rank (the per-step host work - replay bookkeeping, ctypes calls - of 8
    replicas otherwise migrates over, and contends for, the same cores).  Returns the core list."""
    try:
        cores = sorted(os.sched_getaffinity(0))
        per = max(1, len(cores) // max(1, world))
        mine = cores[(local_rank * per) % len(cores):][:per] or cores
        os.sched_setaffinity(0, mine)
        return mine
    except (AttributeError, OSError):
        return []


class Timer:
    """Device timing shared by every leg: barrier + synchronize on both sides, CUDA events on the launching
    stream, max over ranks.  windows > 1: `windows` consecutive windows of `steps` calls, an event between
    them, the median window (max over ranks per window) is the result."""

    def __init__(self, torch, dist):
        self.torch, self.dist = torch, dist

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def __call__(self, fn, steps, windows=1):
        torch = self.torch
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(windows + 1)]
        self.barrier()
        ev[0].record()
        for w in range(windows):
            for _ in range(steps):
                fn()
            ev[w + 1].record()
        self.barrier()
        ms = torch.tensor([ev[w].elapsed_time(ev[w + 1]) for w in range(windows)], device="cuda", dtype=torch.float64)
        if self.dist is not None:
            self.dist.all_reduce(ms, op=self.dist.ReduceOp.MAX)
        ms = ms.cpu().numpy()
        self.last = {"count": int(windows), "median_ms": float(np.median(ms)), "min_ms": float(ms.min()),
                     "max_ms": float(ms.max()), "first_ms": float(ms[0])}
        return float(np.median(ms))


def run_ours(args, rank, world, local_rank):
    import torch
    from dqn_mgsc_zoo_b200 import learner as lrn
    from dqn_mgsc_zoo_b200 import replay as rp
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local_rank)
    cores = pin_to_cores(local_rank, world)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    timed = Timer(torch, dist)

    # ---- build the workload: full replay + learner
    t_fill = time.perf_counter()
    replay = rp.TransitionReplay(args.capacity, rp.Transition(None, None, None, None, None),
                                 np.random.RandomState(rank), device=local_rank)
    src = EpisodeSource(1000 + rank)
    # a sparse host-side record of what went in, for the bit-exact check at the benchmarked scale
    keep_ids = set(int(i) for i in np.random.default_rng(7).choice(args.capacity, size=min(64, args.capacity),
                                                                  replace=False))
    keep_ids |= {0, args.capacity - 1}
    kept = {}
    for i in range(args.capacity):
        tr = src.next()
        replay.add(rp.Transition(*tr))
        if i in keep_ids:
            kept[i] = tr
    store = replay.frame_store
    store.flush()
    t_fill = time.perf_counter() - t_fill

    def check_scale(ids):
        """gather (slot = id mod capacity) == the host transitions that were added, byte for byte."""
        ids = sorted(ids)
        got = store.gather_host(np.asarray(ids, dtype=np.int64) % args.capacity)
        ok = True
        for row, i in enumerate(ids):
            s0, a, r, d, s1 = kept[i]
            ok &= bool(np.array_equal(got[0][row], s0) and np.array_equal(got[4][row], s1) and got[1][row] == a
                       and got[2][row] == np.float32(r) and got[3][row] == np.float32(d))
        return ok

    parity_at_scale = check_scale(kept.keys())

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

    for _ in range(max(3, args.warmup)):
        device_step()
    windows = max(1, args.windows)
    with ClockSampler(local_rank) as clocks:
        ms = timed(device_step, args.steps, windows)
    value = world * args.steps / (ms * 1e-3)
    win_value = dict(timed.last)

    # ---- e2e: new transitions arrive from host memory every step, loss is read back
    e2e_windows = max(1, min(windows, 12000 // max(1, args.steps)))
    e2e_total = args.steps * e2e_windows
    sync_steps = min(args.steps, 300)
    # the e2e leg is paced by the HOST (Python replay bookkeeping + driver calls ~60 us per step against a ~64 us
    # device step): its warm-up is longer than --warmup so that the interpreter, the CPU caches over the 1 M-entry
    # id tables and the host cores' clocks are in steady state before the timed windows (untimed, stated in the line)
    e2e_warm = max(300, args.warmup)
    pending = collections.deque(src.next() for _ in range(ADDS_PER_LEARN * (e2e_total + sync_steps + e2e_warm + 10)))
    for _ in range(e2e_warm):
        e2e_step()
    up0 = store.stats()["planes_uploaded"]
    ms_e2e = timed(e2e_step, args.steps, e2e_windows)     # the closing barrier waits for the last step's update
    win_e2e = dict(timed.last)
    last_loss = L.loss_take()                 # drain: the final step's loss
    planes = store.stats()["planes_uploaded"] - up0
    h2d = planes * PLANE / e2e_total + ADDS_PER_LEARN * 64 + BATCH * 8
    for _ in range(3):
        e2e_step_sync()
    ms_sync = timed(e2e_step_sync, sync_steps)
    e2e = {"value": world * args.steps / (ms_e2e * 1e-3), "unit": "updates/s",
           "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": 4, "steps": args.steps, "windows": win_e2e,
           "warmup_steps": e2e_warm,
           "what": "per step: 4 replay.add(host transition) [H2D of the new frames via pinned staging] + "
                   "sample_slots + update_from_replay + D2H of the step's loss; the loss copy is queued "
                   "behind the update and consumed one step later, so host work overlaps the device",
           "blocking_loss_updates_per_s": world * sync_steps / (ms_sync * 1e-3),
           "last_loss": last_loss}
    # the ring wrapped during the e2e legs: the surviving part of the sparse record must still read back exactly
    t_now = replay._t
    parity_at_scale &= check_scale([i for i in kept if i >= t_now - args.capacity])

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
    traffic, traffic_file = ncu_gather_traffic()
    roofline = {"kernel": "gather_stack4", "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                "transitions_per_launch": n, "us_per_launch": ms_g * 1e3,
                "algorithmic_bytes_per_launch": n * GATHER_BYTES_PER_TRANSITION,
                "note": "CUDA events around %d back-to-back launches on the launching stream, slot lists already on "
                        "the device; 24 different random slot lists over a %.1f GB plane pool (>> L2). traffic = "
                        "dram read+write per launch of the same kernel and n from profiles/%s (ncu --set full of "
                        "scripts/ncu_target.py: a 60k-transition pool, not this run's; the write-back of the last "
                        "lines is still in L2 when the kernel ends)"
                        % (greps, store.stats()["hbm_bytes"] / 1e9, traffic_file)}
    ms_upd = ms / args.steps
    roofline_update = {"bound": "hbm", "achieved": UPDATE_BYTES / (ms_upd * 1e-3) / 1e9, "peak": peak,
                       "unit": "GB/s", "frac": UPDATE_BYTES / (ms_upd * 1e-3) / 1e9 / peak,
                       "note": "whole update graph vs 49.05 MB touched-once bytes; batch 32 is latency bound"}

    # ---- the same step with the GEMMs in plain fp32 (SIMT path), for the record
    other = {}
    if args.precision == "tf32" and not args.no_extras:
        L32 = lrn.Learner(NUM_ACTIONS, BATCH, shared_bias=True, double_q=True, precision="fp32", device=local_rank)
        L32.set_params(init, lrn.ONLINE)
        L32.set_params(init, lrn.TARGET)

        def fp32_step():
            L32.update_from_replay(store, replay.sample_slots(BATCH))
        for _ in range(10):
            fp32_step()
        n32 = min(args.steps, 300)
        other["fp32_simt_updates_per_s"] = world * n32 / (timed(fp32_step, n32) * 1e-3)
        L32.close()

    launches = L.launches_per_update() + 1
    replay_stats = store.stats()
    extras = {}
    if not args.no_extras:
        # free the headline workload first: the sub-records build their own stores
        del slot_sets, bufs
        L.close()
        replay = store = None
        import gc
        gc.collect()
        torch.cuda.empty_cache()
        extras["agent_loop"] = bench_agent_loop(args, local_rank, rank)
        extras["reservoir"] = bench_reservoir(args, timed, local_rank, rank)
        extras["mgsc"] = bench_mgsc(args, timed, local_rank, rank)
        extras["dp"] = bench_dp_record(args, timed, local_rank, rank, world, dist)

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        res = cpu_reference_arm(args.cpu_seconds, capacity=args.cpu_capacity)
        cpu = {k: res[k] for k in ("value", "unit", "cores", "kind", "sample")}
    line = {
        "metric": "learner_updates_per_s", "value": value, "unit": "updates/s", "n_gpus": world,
        "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32" if args.precision == "fp32" else args.precision, "data": "synthetic",
        "config": workload_config(args), "clocks": clocks.summary(),
        "timing": "median of %d consecutive windows of `steps` updates (barrier + synchronize before the first and "
                  "after the last window, one CUDA event between windows, max over ranks per window)" % windows,
        "windows": win_value, "host_cores_per_rank": len(cores), "e2e": e2e,
        "gpu_launches": int(launches * args.steps), "launches_per_update": int(launches),
        "roofline": roofline, "roofline_update": roofline_update, "cpu_baseline": cpu,
        "parity_at_scale": bool(parity_at_scale),
        "other_precisions": other, "replay_gather_gbs": achieved, "replay_fill_seconds": t_fill,
        "replay_stats": replay_stats,
    }
    line.update(extras)
    emit(line)
    if dist is not None:
        dist.destroy_process_group()


def bench_agent_loop(args, device, rank):
    """C1 stand-in through the reference's own seams: the DoubleDqn drop-in driven by parts.run_loop
    (agent.step -> preprocess, act, accumulate, replay.add, _learn every 16 frames, target sync) on the synthetic
    environment (ALE is absent and outside the optimisation target).  Host wall clock: this loop is host-paced."""
    from dqn_mgsc_zoo_b200 import agent as agent_lib
    from dqn_mgsc_zoo_b200 import networks, parts, synthetic
    from dqn_mgsc_zoo_b200 import replay as rp
    cap, frames = 50_000, 24_000
    replay = rp.TransitionReplay(cap, rp.Transition(None, None, None, None, None), np.random.RandomState(rank),
                                 device=device)
    ag = agent_lib.DoubleDqn(
        preprocessor=synthetic.StackingPreprocessor(num_action_repeats=4),
        sample_network_input=np.zeros((84, 84, 4), np.uint8), network=networks.double_dqn_atari_network(NUM_ACTIONS),
        optimizer=networks.rmsprop(learning_rate=0.00025, decay=0.95, eps=0.01 / 32 ** 2, centered=True),
        transition_accumulator=rp.TransitionAccumulator(), replay=replay, batch_size=BATCH,
        exploration_epsilon=lambda t: 0.1, min_replay_capacity_fraction=0.02, learn_period=16,
        target_network_update_period=30_000, grad_error_bound=1.0 / 32, rng_key=1 + rank, device=device,
        precision=args.precision)
    env = synthetic.SyntheticEnvironment(np.random.default_rng(3 + rank), NUM_ACTIONS, min_len=50, max_len=400)
    warm = int(0.02 * cap) * 4 + 2000
    n, t0 = 0, None
    for _ in parts.run_loop(ag, env, max_steps_per_episode=27_000):
        n += 1
        if n == warm:
            ag.learner.synchronize()
            t0 = time.perf_counter()
        if n == warm + frames:
            break
    ag.learner.synchronize()
    dt = time.perf_counter() - t0
    out = {"frames_per_s": frames / dt, "learner_updates_per_s": frames / 16 / dt, "frames": frames,
           "precision": ag.precision,
           "what": "agent.DoubleDqn through parts.run_loop on the synthetic environment: action repeat 4, learn every "
                   "16 frames at batch 32, epsilon 0.1, %d-transition replay; host wall clock, one rank (the "
                   "reference's V100 runs: 1 084 train frames/s with ALE, results/dqn/seed_*.csv)" % cap}
    ag.learner.close()
    return out


def bench_reservoir(args, timed, device, rank):
    """C3: reservoir (Algorithm R) insert + sample stress at the headline capacity: fill, then adds with one
    sample(32)+gather every 4 adds.  Host RandomState draws as in replay.py:267-283, slots overwritten in HBM."""
    import torch
    from dqn_mgsc_zoo_b200 import replay as rp
    cap = args.capacity
    res = rp.ReservoirTransitionReplay(cap, rp.Transition(None, None, None, None, None),
                                       np.random.RandomState(50 + rank), device=device)
    src = EpisodeSource(2000 + rank)
    for _ in range(cap):
        res.add(rp.Transition(*src.next()))
    store = res.frame_store
    store.flush()
    n_add = 20000
    trs = [rp.Transition(*src.next()) for _ in range(n_add)]
    bufs = [torch.empty((BATCH, 84, 84, 4), dtype=torch.uint8, device="cuda") for _ in range(2)]
    ai = torch.empty(BATCH, dtype=torch.int32, device="cuda")
    rf, df = torch.empty(BATCH, device="cuda"), torch.empty(BATCH, device="cuda")
    it = iter(trs)
    puts0 = store.stats()["puts"]

    def add4_sample():
        for _ in range(ADDS_PER_LEARN):
            res.add(next(it))
        store.gather(res.sample_slots(BATCH), bufs[0].data_ptr(), bufs[1].data_ptr(), ai.data_ptr(), rf.data_ptr(),
                     df.data_ptr())

    steps = n_add // ADDS_PER_LEARN - 8
    for _ in range(8):
        add4_sample()
    ms = timed(add4_sample, steps)
    st = store.stats()
    out = {"capacity": cap, "adds_per_s": steps * ADDS_PER_LEARN / (ms * 1e-3), "samples_per_s": steps / (ms * 1e-3),
           "accepted_fraction": (st["puts"] - puts0) / float(n_add),
           "what": "%d adds (each: RandomState.randint(0, t) -> replace slot j in HBM or drop) interleaved with one "
                   "sample(32) + gather every 4 adds; device-timed, max over ranks" % (steps * ADDS_PER_LEARN)}
    del res, store
    return out


def bench_mgsc(args, timed, device, rank):
    """C4: one MGSC agent step (dqn_mgsc_batched/agent.py:237-280) phase by phase, M = 5 (the reference's timed
    configuration, run_mgscdqnbatched_normal_timing.sh:50), host-synchronous per phase like the reference's table."""
    import torch
    from dqn_mgsc_zoo_b200 import learner as lrn
    from dqn_mgsc_zoo_b200 import networks
    from dqn_mgsc_zoo_b200 import replay as rp
    from dqn_mgsc_zoo_b200 import replay_circular as rc
    cap, M, B, A = args.capacity, 5, BATCH, NUM_ACTIONS
    pc = time.perf_counter
    src = EpisodeSource(3000 + rank)
    replay = rc.MGSCFiFoTransitionReplay(cap, rp.Transition(None, None, None, None, None),
                                         np.random.default_rng(rank), device=device)
    t0 = pc()
    for _ in range(cap):
        replay.add(rp.Transition(*src.next()))
    torch.cuda.synchronize()
    fill = pc() - t0
    store = replay.frame_store
    L = lrn.Learner(A, B, shared_bias=False, double_q=False, precision=args.precision, device=device)
    init = networks.init_params(np.random.default_rng(0), A, False)
    L.set_params(init, lrn.ONLINE)
    L.set_params(init, lrn.TARGET)
    for _ in range(10):
        L.update_from_replay(store, replay.sample_slots(B))
    meta = lrn.MetaLearner(L, M)
    names = ["replay-add", "replay-sample-meta-batch", "meta-update", "replay-update", "replay-sample-batch", "update"]
    acc = dict.fromkeys(names, 0.0)
    steps = 200
    t_all = 0.0
    for it in range(steps + 5):
        if it == 5:
            acc = dict.fromkeys(names, 0.0)
            torch.cuda.synchronize()
            t_all = pc()
        tr = rp.Transition(*src.next())
        t = pc(); idx, slots, lg = replay.batch_of_ids_slots_and_logits(M); torch.cuda.synchronize()
        acc["replay-sample-meta-batch"] += pc() - t
        t = pc(); new, _ = meta.meta_update_from_replay(store, slots, lg, tr); torch.cuda.synchronize()
        acc["meta-update"] += pc() - t
        t = pc(); replay.update_priorities(idx, new); torch.cuda.synchronize()
        acc["replay-update"] += pc() - t
        t = pc()
        for _ in range(ADDS_PER_LEARN):
            replay.add(rp.Transition(*src.next()))
        torch.cuda.synchronize()
        acc["replay-add"] += (pc() - t) / ADDS_PER_LEARN
        t = pc(); sl = replay.sample_slots(B); torch.cuda.synchronize()
        acc["replay-sample-batch"] += pc() - t
        t = pc(); L.update_from_replay(store, sl); torch.cuda.synchronize()
        acc["update"] += pc() - t
    torch.cuda.synchronize()
    t_all = pc() - t_all
    us = {k: v / steps * 1e6 for k, v in acc.items()}
    # the second-order objective of dqn_mgsc_batched_reservoir (Hessian-vector product through the CNN), same meta batch
    meta2 = lrn.MetaLearner(L, M, second_order=True)
    tr = rp.Transition(*src.next())
    idx, slots, lg = replay.batch_of_ids_slots_and_logits(M)
    for _ in range(3):
        meta2.meta_update_from_replay(store, slots, lg, tr)
    torch.cuda.synchronize()
    t = pc()
    for _ in range(50):
        meta2.meta_update_from_replay(store, slots, lg, tr)
    torch.cuda.synchronize()
    us["meta-update-second-order"] = (pc() - t) / 50 * 1e6
    meta2.close()
    out = {"capacity": cap, "meta_batch": M, "phase_us": us, "step_us": t_all / steps * 1e6,
           "steps_per_s": steps / t_all, "fill_seconds": fill,
           "reference_phase_us_v100": {"replay-add": 800, "replay-sample-meta-batch": 55600, "meta-update": 4100,
                                       "replay-update": 540, "replay-sample-batch": 57100, "update": 12200},
           "what": "per step: meta batch of %d uniform ids + logits, meta_update, update_priorities, 4 adds "
                   "(logit = lse(all) - log(size) each), logit-weighted sample(32), update; host-synchronous per phase; "
                   "reference column: dqn_mgsc_batched/run_atari.py:308-316 (V100, its own dispatch-timed table)" % M}
    meta.close()
    L.close()
    del replay, store
    return out


def bench_dp_record(args, timed, device, rank, world, dist):
    """C5-ii next to the replica curve: ONE data-parallel learner, global batch 4096 split over the ranks,
    gradient all-reduce (NCCL) in the library; strong scaling."""
    from dqn_mgsc_zoo_b200 import learner as lrn
    from dqn_mgsc_zoo_b200 import networks, parallel
    from dqn_mgsc_zoo_b200 import replay as rp
    import torch
    G = args.global_batch
    if G % world != 0:
        return {"skipped": "global batch %d does not split over %d ranks" % (G, world)}
    capacity = min(args.capacity, 50_000)      # replicated ring; the sampling pattern, not the size, matters here
    replay = rp.TransitionReplay(capacity, rp.Transition(None, None, None, None, None),
                                 np.random.RandomState(0), device=device)      # same seed on every rank: replicas
    src = EpisodeSource(1000)
    for _ in range(capacity):
        replay.add(rp.Transition(*src.next()))
    store = replay.frame_store
    store.flush()
    dp = parallel.make_nccl_learner(NUM_ACTIONS, G, device, rank, world, shared_bias=True, double_q=True,
                                    precision=args.precision)
    L = dp.learner
    init = networks.init_params(np.random.default_rng(0), NUM_ACTIONS, shared_bias=True)
    L.set_params(init, lrn.ONLINE)
    L.set_params(init, lrn.TARGET)

    def step():
        dp.update_from_replay(store, replay.sample_slots(G))      # identical draw on every rank

    for _ in range(5):
        step()
    steps = 40
    ms = timed(step, steps, 5)
    ar_us = None
    if world > 1:
        dp.all_reduce_only()
        ar_us = timed(dp.all_reduce_only, 50) / 50 * 1e3
    digest = float(np.abs(L.read(lrn.ONLINE).astype(np.float64)).sum())
    same = True
    if world > 1:
        t = torch.tensor([digest], device="cuda", dtype=torch.float64)
        lo, hi = t.clone(), t.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        same = bool(lo.item() == hi.item())
    value = steps / (ms * 1e-3)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops", 1398.4))
    tf = UPDATE_FLOP_PER_SAMPLE * G * value / 1e12 / world
    out = {"global_batch": G, "rows_per_gpu": G // world, "updates_per_s": value, "samples_per_s": value * G,
           "ms_per_step": ms / steps, "scaling": "strong", "allreduce_us": ar_us, "replicas_identical": same,
           "tflops_per_gpu": tf, "frac_of_bf16_peak": tf / peak, "collective": dp.describe(),
           "what": "global batch %d, %d rows per GPU, replicated %d-transition replay, identical index draw on every "
                   "rank; forward/TD/backward on the local rows with the global loss mean, all-reduce(sum) of the "
                   "6.75 MB gradient arena, identical RMSProp step everywhere; median of 5 windows of %d updates"
                   % (G, G // world, capacity, steps)}
    dp.close()
    del replay, store
    return out


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
