"""Short program for `ncu` on the large-batch (stream-kernel) learner: two tf32 updates at batch B
(argv[1], default 2048) from host batches.  Graph replay is profiled kernel by kernel."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from dqn_mgsc_zoo_b200 import learner as lrn, networks, synthetic
B = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
L = lrn.Learner(6, B, shared_bias=True, double_q=True, precision="tf32")
init = networks.init_params(np.random.default_rng(0), 6, True)
L.set_params(init, lrn.ONLINE); L.set_params(init, lrn.TARGET)
batch = synthetic.random_batch(np.random.default_rng(1), B, 6)
for _ in range(2):
    L.update_host(*batch)
L.synchronize()
print("ok", L.loss())
