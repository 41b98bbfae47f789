"""Debug aid: per-step drift of the CUDA learner vs the float64 / float32 oracles."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from dqn_mgsc_zoo_b200 import learner as lrn, synthetic
from oracle import learner_oracle as net, update_oracle as upd

def rel(x, ref):
    x, ref = np.asarray(x, np.float64), np.asarray(ref, np.float64)
    return np.abs(x - ref).max() / (np.abs(ref).max() + 1e-30)

B, A, steps = 32, 6, int(sys.argv[1]) if len(sys.argv) > 1 else 30
rng = np.random.default_rng(2)
online = net.init_params(rng, A, False); target = net.init_params(rng, A, False)
L = lrn.Learner(A, B); L.set_params(online, lrn.ONLINE); L.set_params(target, lrn.TARGET)
P = L.num_params
o64 = dict(p=online, t=target, mu=np.zeros(P), nu=np.zeros(P))
o32 = dict(p={k: v.astype(np.float32) for k, v in online.items()}, t={k: v.astype(np.float32) for k, v in target.items()},
           mu=np.zeros(P, np.float32), nu=np.zeros(P, np.float32))
h32 = upd.RmsProp(np.float32(0.00025), np.float32(0.95), np.float32(0.01 / 1024))
keys = list(online)
for step in range(steps):
    batch = synthetic.random_batch(rng, B, A)
    outs = []
    for o, h in ((o64, upd.DQN_RMSPROP), (o32, h32)):
        out = upd.update(o["p"], o["t"], o["mu"], o["nu"], batch, h=h)
        o["p"], o["mu"], o["nu"] = out["params"], out["mu"], out["nu"]
        outs.append(out)
    L.update_host(*batch)
    g = lrn.unflatten_params(L.read(lrn.GRAD), A, False)
    p = L.get_params()
    line = "step %2d loss %.5f/%.5f |" % (step, L.loss(), outs[0]["loss"])
    for k in ("conv1.w", "conv2.w", "fc1.w", "fc2.w"):
        line += " %s g %.1e/%.1e p %.1e/%.1e |" % (k[:5], rel(g[k], outs[0]["grads"][k]), rel(outs[1]["grads"][k], outs[0]["grads"][k]),
                                                   rel(p[k], o64["p"][k]), rel(o32["p"][k], o64["p"][k]))
    print(line, "mu %.1e nu %.1e" % (rel(L.read(lrn.MU), o64["mu"]), rel(L.read(lrn.NU), o64["nu"])), flush=True)
    if step in (0, 5, 20):
        # where is the worst param element and what do its g, mu, nu look like
        k = "conv1.w"
        d = np.abs(p[k].astype(np.float64) - o64["p"][k]); i = np.unravel_index(d.argmax(), d.shape)
        fl = np.ravel_multi_index(i, d.shape)
        print("   worst conv1.w elem", i, "gpu p %.9f ref %.9f f32 %.9f | g gpu %.3e ref %.3e | mu %.3e/%.3e nu %.3e/%.3e" % (
            p[k][i], o64["p"][k][i], o32["p"][k][i], g[k][i], outs[0]["grads"][k][i],
            L.read(lrn.MU)[fl], o64["mu"][fl], L.read(lrn.NU)[fl], o64["nu"][fl]))
