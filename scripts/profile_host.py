"""Host-side cost of one e2e step (GPU box): replay.add, sample_slots, update call, loss readback.
   python scripts/profile_host.py [precision]"""
import collections, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench
from dqn_mgsc_zoo_b200 import learner as lrn, networks, replay as rp

precision = sys.argv[1] if len(sys.argv) > 1 else "tf32"
CAP = int(sys.argv[2]) if len(sys.argv) > 2 else 100_000
replay = rp.TransitionReplay(CAP, rp.Transition(None, None, None, None, None), np.random.RandomState(0), device=0)
src = bench.EpisodeSource(1000)
for _ in range(CAP):
    replay.add(rp.Transition(*src.next()))
store = replay.frame_store
store.flush()
L = lrn.Learner(bench.NUM_ACTIONS, bench.BATCH, shared_bias=True, double_q=True, precision=precision)
init = networks.init_params(np.random.default_rng(0), bench.NUM_ACTIONS, shared_bias=True)
L.set_params(init, lrn.ONLINE); L.set_params(init, lrn.TARGET)
N = 2000
pending = [rp.Transition(*src.next()) for _ in range(10 * N)]
for _ in range(20):
    L.update_from_replay(store, replay.sample_slots(32))
torch.cuda.synchronize()

def clock(name, fn, n, sync=True):
    if sync: torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    if sync: torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / n * 1e6
    print("%-44s %8.2f us" % (name, dt))
    return dt

it = iter(pending)
clock("replay.add (host transition)", lambda: replay.add(next(it)), 4 * N // 2)
clock("frame_store.put only", lambda: store.put(5, *[pending[7][k] for k in (0, 4, 1, 2, 3)]), 2000)
clock("sample_slots(32)", lambda: replay.sample_slots(32), N)
slots = replay.sample_slots(32)
clock("update_from_replay (device-bound loop)", lambda: L.update_from_replay(store, slots), N)
def upd_sync():
    L.update_from_replay(store, slots); return L.loss()
clock("update_from_replay + loss() sync", upd_sync, N)
clock("loss() alone (idle device)", L.loss, N)
def full():
    for _ in range(4):
        replay.add(next(it))
    L.update_from_replay(store, replay.sample_slots(32))
    return L.loss()
clock("full e2e step (sync)", full, N // 2)
if hasattr(L, "loss_lagged"):
    def full_lag():
        for _ in range(4):
            replay.add(next(it))
        L.update_from_replay(store, replay.sample_slots(32))
        return L.loss_lagged()
    clock("full e2e step (loss one step behind)", full_lag, N // 2 - 10)

# host-time breakdown of the pipelined step (device runs asynchronously underneath)
acc = collections.OrderedDict((k, 0.0) for k in ("add x4", "sample_slots", "update_from_replay", "loss_lagged"))
pc = time.perf_counter
torch.cuda.synchronize()
n = 1000
t_all = pc()
for _ in range(n):
    t0 = pc()
    for _ in range(4):
        replay.add(next(it))
    t1 = pc()
    sl = replay.sample_slots(32)
    t2 = pc()
    L.update_from_replay(store, sl)
    t3 = pc()
    L.loss_lagged()
    t4 = pc()
    acc["add x4"] += t1 - t0; acc["sample_slots"] += t2 - t1; acc["update_from_replay"] += t3 - t2
    acc["loss_lagged"] += t4 - t3
torch.cuda.synchronize()
t_all = pc() - t_all
for k, v in acc.items():
    print("  host %-22s %8.2f us/step" % (k, v / n * 1e6))
print("  total %.2f us/step" % (t_all / n * 1e6))
