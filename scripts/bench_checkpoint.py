"""N2 at the benchmarked scale: get_state -> set_state round trip of a 1 M-transition HBM replay
(dqnb_replay_export / import), time and identical subsequent samples.
   python scripts/bench_checkpoint.py [capacity]"""
import copy, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from dqn_mgsc_zoo_b200 import replay as rp

cap = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
structure = rp.Transition(None, None, None, None, None)
p = rp.TransitionReplay(cap, structure, np.random.RandomState(0))
src = bench.EpisodeSource(11)
t0 = time.perf_counter()
for _ in range(cap + 1000):
    p.add(rp.Transition(*src.next()))
t_fill = time.perf_counter() - t0
t0 = time.perf_counter()
state = p.get_state()
t_get = time.perf_counter() - t0
blob_gb = state['storage'].snapshot.blob.nbytes / 1e9
# no pickle in between here: un-alias the host tables the two replays would otherwise share (a checkpoint file does this)
state = dict(state, distribution=copy.deepcopy(state['distribution']))
q = rp.TransitionReplay(cap, structure, np.random.RandomState(1))
t0 = time.perf_counter()
q.set_state(state)
t_set = time.perf_counter() - t0
p._random_state.seed(5); q._random_state.seed(5)
same = True
for _ in range(4):
    a, b = p.sample(64), q.sample(64)
    same &= all(np.array_equal(x, y) for x, y in zip(a, b))
for _ in range(200):                      # both keep accepting adds identically
    tr = rp.Transition(*src.next())
    p.add(tr); q.add(tr)
a, b = p.sample(64), q.sample(64)
same &= all(np.array_equal(x, y) for x, y in zip(a, b))
i0, item0 = state['storage'][12345]
want = next(p.get([i0])) if i0 in p._live else None
print(json.dumps({"capacity": cap, "fill_s": t_fill, "get_state_s": t_get, "set_state_s": t_set,
                  "round_trip_s": t_get + t_set, "snapshot_gb": blob_gb, "identical_samples_after": bool(same),
                  "lazy_item_matches": bool(want is None or all(np.array_equal(x, y) for x, y in zip(item0, want)))}))
