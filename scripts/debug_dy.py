import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from dqn_mgsc_zoo_b200 import learner as lrn, synthetic
from oracle import learner_oracle as net, update_oracle as upd
def rel(x, ref):
    x, ref = np.asarray(x, np.float64), np.asarray(ref, np.float64)
    return np.abs(x - ref).max() / (np.abs(ref).max() + 1e-30)
B, A = 32, 6
for seed in range(6):
    rng = np.random.default_rng(seed)
    online = net.init_params(rng, A, False); target = net.init_params(rng, A, False)
    L = lrn.Learner(A, B); L.set_params(online, lrn.ONLINE); L.set_params(target, lrn.TARGET)
    batch = synthetic.random_batch(rng, B, A)
    out = upd.loss_and_grads(online, target, batch)
    L.update_host(*batch)
    c = out["cache"]
    g = lrn.unflatten_params(L.read(lrn.GRAD), A, False)
    line = "seed %d" % seed
    for name, wid, aid in (("fc1", 100, None), ("conv3", 101, 111), ("conv2", 102, 112), ("conv1", 103, 113)):
        dy = L.read(wid).reshape(c["dy." + name].shape)
        line += " | %s dy %.1e" % (name, rel(dy, c["dy." + name]))
        if aid:
            act = L.read(aid).reshape(c[name + ".out"].shape)
            flips = int(((act > 0) != (c[name + ".out"] > 0)).sum())
            line += " act %.1e flips %d" % (rel(act, c[name + ".out"]), flips)
            bad = np.abs(dy - c["dy." + name]) > 1e-4 * np.abs(c["dy." + name]).max()
            line += " bad %d" % int(bad.sum())
            if bad.sum():
                idx = np.argwhere(bad)[:3]
                line += " at %s" % idx.tolist()
        line += " gw %.1e gb %.1e" % (rel(g[name + ".w"], out["grads"][name + ".w"]), rel(g[name + ".b"], out["grads"][name + ".b"]))
    print(line, flush=True)
