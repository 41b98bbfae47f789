"""Per-kernel timing of the learner update on the GPU (CUDA events, warm caches, real clocks).
   python scripts/profile_update.py [precision] [batch]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from dqn_mgsc_zoo_b200 import learner as lrn, networks, synthetic

precision = sys.argv[1] if len(sys.argv) > 1 else "tf32"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 32
A = 6
L = lrn.Learner(A, B, shared_bias=True, double_q=True, precision=precision)
init = networks.init_params(np.random.default_rng(0), A, True)
L.set_params(init, lrn.ONLINE)
L.set_params(init, lrn.TARGET)
batch = synthetic.random_batch(np.random.default_rng(1), B, A)
for _ in range(5):
    L.update_host(*batch)
L.synchronize()
print("precision %s batch %d: %d launches per update" % (precision, B, L.launches_per_update()))
print(L.profile(50))
print(L.timeline())
import torch, time
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s1, a, r, d, s2 = batch
import ctypes
from dqn_mgsc_zoo_b200 import _lib
# graph replay only (batch already in the learner's buffers): device time per update
lib = _lib.load()
n = 300
torch.cuda.synchronize()
ev0.record()
for _ in range(n):
    L.update_host(*batch)
ev1.record()
torch.cuda.synchronize()
print("update_host loop: %.1f us/update" % (ev0.elapsed_time(ev1) * 1e3 / n))

# graph replays only (batch resident, no copies, no gather): period of back-to-back updates
torch.cuda.synchronize()
_lib.call("dqnb_learner_relaunch", L._h, 50, L.stream)
torch.cuda.synchronize()
ev0.record()
_lib.call("dqnb_learner_relaunch", L._h, 1000, L.stream)
ev1.record()
torch.cuda.synchronize()
print("graph replay loop: %.2f us/update" % (ev0.elapsed_time(ev1)))
