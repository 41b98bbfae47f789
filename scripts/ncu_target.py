"""Short program for `ncu --set full`: the gather kernel at 4096 transitions per launch (the
HBM-roofline configuration of bench.py) three times, then two tf32 updates from the replay."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from dqn_mgsc_zoo_b200 import learner as lrn, networks, replay as rp
cap = 60000
replay = rp.TransitionReplay(cap, rp.Transition(None, None, None, None, None), np.random.RandomState(0))
src = bench.EpisodeSource(5)
for _ in range(cap):
    replay.add(rp.Transition(*src.next()))
store = replay.frame_store
store.flush()
L = lrn.Learner(6, 32, shared_bias=True, double_q=True, precision="tf32")
init = networks.init_params(np.random.default_rng(0), 6, True)
L.set_params(init, lrn.ONLINE); L.set_params(init, lrn.TARGET)
n = 4096
bufs = [torch.empty((n, 84, 84, 4), dtype=torch.uint8, device="cuda") for _ in range(2)]
ai = torch.empty(n, dtype=torch.int32, device="cuda"); rf = torch.empty(n, device="cuda"); df = torch.empty(n, device="cuda")
rs = np.random.RandomState(1)
for _ in range(3):
    store.gather(rs.randint(cap, size=n).astype(np.int64), bufs[0].data_ptr(), bufs[1].data_ptr(), ai.data_ptr(),
                 rf.data_ptr(), df.data_ptr())
for _ in range(2):
    L.update_from_replay(store, replay.sample_slots(32))
obs = np.random.default_rng(2).integers(0, 256, (84, 84, 4), dtype=np.uint8)
for _ in range(2):
    q = L.q_values(obs)        # acting graph: act_forward_kernel
torch.cuda.synchronize()
print("ok", L.loss(), q)
