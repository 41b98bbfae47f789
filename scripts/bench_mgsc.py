"""C4: per-phase time of one MGSC agent step at replay capacity `cap`, on the device path.
Phases named like the reference's own timing table (dqn_mgsc_batched/agent.py:84-91):
replay-add, replay-sample-meta-batch, meta-update, replay-update, replay-sample-batch, update.
   python scripts/bench_mgsc.py [cap] [precision] [M ...]
Also C3: reservoir insert+sample stress (adds/s, samples/s)."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench
from dqn_mgsc_zoo_b200 import learner as lrn, networks, replay as rp, replay_circular as rc

cap = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
precision = sys.argv[2] if len(sys.argv) > 2 else "fp32"
Ms = [int(x) for x in sys.argv[3:]] or [5, 10, 100]
A, B = bench.NUM_ACTIONS, bench.BATCH
pc = time.perf_counter
out = {"capacity": cap, "precision": precision}

src = bench.EpisodeSource(7)
replay = rc.MGSCFiFoTransitionReplay(cap, rp.Transition(None, None, None, None, None), np.random.default_rng(0))
t0 = pc()
for _ in range(cap):
    replay.add(rp.Transition(*src.next()))
torch.cuda.synchronize()
out["fill_seconds"] = pc() - t0
store = replay.frame_store
L = lrn.Learner(A, B, shared_bias=False, double_q=False, precision=precision)
init = networks.init_params(np.random.default_rng(0), A, False)
L.set_params(init, lrn.ONLINE); L.set_params(init, lrn.TARGET)
for _ in range(20):
    L.update_from_replay(store, replay.sample_slots(B))
torch.cuda.synchronize()

def phase_times(M, steps):
    meta = lrn.MetaLearner(L, M)
    acc = dict.fromkeys(["replay-add", "replay-sample-meta-batch", "meta-update", "replay-update",
                         "replay-sample-batch", "update"], 0.0)
    for it in range(steps + 3):
        if it == 3:
            acc = dict.fromkeys(acc, 0.0)
        tr = rp.Transition(*src.next())
        t = pc(); idx, slots, lg = replay.batch_of_ids_slots_and_logits(M); torch.cuda.synchronize()
        acc["replay-sample-meta-batch"] += pc() - t
        t = pc(); new, loss = meta.meta_update_from_replay(store, slots, lg, tr); torch.cuda.synchronize()
        acc["meta-update"] += pc() - t
        t = pc(); replay.update_priorities(idx, new); torch.cuda.synchronize()
        acc["replay-update"] += pc() - t
        t = pc()
        for _ in range(bench.ADDS_PER_LEARN):
            replay.add(rp.Transition(*src.next()))
        torch.cuda.synchronize()
        acc["replay-add"] += (pc() - t) / bench.ADDS_PER_LEARN
        t = pc(); sl = replay.sample_slots(B); torch.cuda.synchronize()
        acc["replay-sample-batch"] += pc() - t
        t = pc(); L.update_from_replay(store, sl); torch.cuda.synchronize()
        acc["update"] += pc() - t
    meta.close()
    res = {k: v / steps * 1e6 for k, v in acc.items()}
    res["step_us"] = sum(v if k != "replay-add" else v * bench.ADDS_PER_LEARN for k, v in res.items())
    return res

out["mgsc_phase_us"] = {str(M): phase_times(M, 200 if M <= 10 else 40) for M in Ms}
del replay

# C3: reservoir stress at the same capacity: fill, then adds interleaved with sample(32) every 4 adds
res = rp.ReservoirTransitionReplay(cap, rp.Transition(None, None, None, None, None), np.random.RandomState(0))
for _ in range(cap):
    res.add(rp.Transition(*src.next()))
torch.cuda.synchronize()
n_add = 100_000
trs = [rp.Transition(*src.next()) for _ in range(n_add)]
t_add = t_s = 0.0
for i, tr in enumerate(trs):
    t = pc(); res.add(tr); t_add += pc() - t
    if i % 4 == 3:
        t = pc(); L.update_from_replay(res.frame_store, res.sample_slots(B)); t_s += pc() - t
torch.cuda.synchronize()
st = res.frame_store.stats()
out["reservoir"] = {"adds_per_s": n_add / t_add, "sample_plus_update_calls_per_s": (n_add // 4) / t_s,
                    "accepted_fraction_expected": "N/_t ~ 1.0 -> 0.91 over the run", "store": st}
print(json.dumps(out))
