"""Oracle backward conditioned on the DEVICE's own forward (shared by the learner / MGSC GPU tests).

A reduced-precision (tf32) forward perturbs pre-activations by ~5e-4 relative; a ReLU unit whose
pre-activation is within that of zero lands on the other side, which is a discontinuity of the function,
not an arithmetic error of the backward kernels.  So the float64 oracle backward is run through the
activations (hence masks) and td errors the device computed: what is compared is the arithmetic of the
dgrad / wgrad / bias kernels.  The RAW (unconditioned) comparison is asserted separately, at a looser bound."""
import numpy as np

from oracle import learner_oracle as net

TAP = {"conv1": 113, "conv2": 112, "conv3": 111, "h1": 110}     # dqnb_learner_read debug taps
SHAPE = {"conv1": (20, 20, 32), "conv2": (9, 9, 64), "conv3": (7, 7, 64)}


def device_cache(read, s_tm1, B):
    """Forward cache of oracle.learner_oracle.backward with the device's activations.  `read(tap)` returns
    the flat float32 array of a debug tap."""
    dt = np.float64
    x0 = s_tm1.astype(dt) / 255.0
    cache = {"x0": x0}
    x = x0
    for name, k, s, _, _ in net.LAYERS:
        p = net._patches(x, k, s)
        cache[name + ".cols"] = p.reshape(p.shape[0] * p.shape[1] * p.shape[2], -1)
        x = read(TAP[name]).reshape((B,) + SHAPE[name]).astype(dt)
        cache[name + ".out"] = x
    cache["h0"] = x.reshape(B, -1)
    cache["h1"] = read(TAP["h1"]).reshape(B, 512).astype(dt)
    return cache


def row_cache(cache, i):
    """The batch-of-1 cache of row i (per-sample gradients)."""
    out = {"x0": cache["x0"][i:i + 1], "h0": cache["h0"][i:i + 1], "h1": cache["h1"][i:i + 1]}
    for name in SHAPE:
        n = SHAPE[name][0] * SHAPE[name][1]
        out[name + ".cols"] = cache[name + ".cols"][i * n:(i + 1) * n]
        out[name + ".out"] = cache[name + ".out"][i:i + 1]
    return out


def grads(online, cache, td, a, denom, grad_bound=1.0 / 32, weights=None, weight_inside=False):
    """Flat float64 gradient of mean-over-`denom` 0.5*td^2 with rlax.clip_gradient on the cotangent, through
    `cache`.  weights: per-row factor outside the clip (MGSC p_i) or inside it (prioritized replay)."""
    B = len(td)
    w = np.ones(B) if weights is None else np.asarray(weights, np.float64)
    cot = np.clip(td * w / denom, -grad_bound, grad_bound) if weight_inside else \
        np.clip(td / denom, -grad_bound, grad_bound) * w
    dq = np.zeros((B, online["fc2.w"].shape[1]))
    dq[np.arange(B), np.asarray(a)] = -cot
    return net.flatten(net.backward({k: v.astype(np.float64) for k, v in online.items()}, dict(cache), dq))
