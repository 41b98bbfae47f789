"""TEST INFRASTRUCTURE ONLY — NumPy restatement of the reference learner math.

PARITY UNPINNED at the jax/haiku/optax/rlax boundary: those packages (jax
0.3.x, dm-haiku 0.0.6, optax 0.1.2, rlax 0.1.5 — requirements_cc.txt) are not
vendored in /root/reference and not installable here, and the reference's own
tests hold no numeric golden vectors for the learner.  This file
restates their published algorithms, follows the reference call sites below and
the op order of the reference's jaxpr dump `Digraph.gv`, and is self-checked in
tests/test_oracle_learner.py against torch-CPU float64 autograd, finite
differences, and the two known-answer fragments the reference tests hold
(networks_test.py:86-103 shared bias; run_mgsc_test.py:198-218 softmax weights).

Follows:
  networks.py:58-79    U(+-1/sqrt(fan_in)) init for w and b
  networks.py:82-134   conv (NHWC/HWIO/VALID), linear, linear_with_shared_bias
  networks.py:181-221  dqn_torso (x/255, 3 conv+relu, flatten h,w,c), value head
  dqn/agent.py:85-117  loss_fn / update (q_learning, clip_gradient, l2_loss, mean)
  double_q/agent.py:85-111  double-Q variant (selector = online net on s_t)
  dqn/run_atari.py:208-213  optax.rmsprop(lr, decay, eps, centered=True)
  dqn_mgsc_batched/agent.py:104-220  MGSC meta loss / meta update (optax.adam)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference
arm may import this module.  Works in float64 (default) or float32.
"""
import collections

import numpy as np

LAYERS = (  # name, kernel hw, stride, in_ch, out_ch
    ("conv1", 8, 4, 4, 32),
    ("conv2", 4, 2, 32, 64),
    ("conv3", 3, 1, 64, 64),
)
FC1_IN, FC1_OUT = 7 * 7 * 64, 512
HAIKU_NAMES = {  # our name -> haiku module path (run_mgsc_test.py:207-213)
    "conv1": "sequential/sequential/conv2_d",
    "conv2": "sequential/sequential/conv2_d_1",
    "conv3": "sequential/sequential/conv2_d_2",
    "fc1": "sequential/sequential_1/linear",
    "fc2": "sequential/sequential_1/linear_1",
}


def param_shapes(num_actions, shared_bias=False):
    """Ordered leaf shapes of the Nature CNN; this order is the flat arena order."""
    shapes = collections.OrderedDict()
    for name, k, _, cin, cout in LAYERS:
        shapes[name + ".w"] = (k, k, cin, cout)
        shapes[name + ".b"] = (cout,)
    shapes["fc1.w"] = (FC1_IN, FC1_OUT)
    shapes["fc1.b"] = (FC1_OUT,)
    shapes["fc2.w"] = (FC1_OUT, num_actions)
    shapes["fc2.b"] = (1,) if shared_bias else (num_actions,)
    return shapes


def init_params(rng, num_actions, shared_bias=False, dtype=np.float64):
    """networks.py:58-79: every leaf ~ U(-c, c), c = 1/sqrt(fan_in) (biases too)."""
    params = collections.OrderedDict()
    for key, shape in param_shapes(num_actions, shared_bias).items():
        layer = key.split(".")[0]
        if layer.startswith("conv"):
            _, k, _, cin, _ = [l for l in LAYERS if l[0] == layer][0]
            fan_in = k * k * cin
        else:
            fan_in = FC1_IN if layer == "fc1" else FC1_OUT
        c = np.sqrt(1.0 / fan_in)
        params[key] = rng.uniform(-c, c, size=shape).astype(dtype)
    return params


def flatten(params):
    return np.concatenate([np.asarray(v).ravel() for v in params.values()])


def unflatten(flat, num_actions, shared_bias=False):
    out, o = collections.OrderedDict(), 0
    for key, shape in param_shapes(num_actions, shared_bias).items():
        n = int(np.prod(shape))
        out[key] = np.asarray(flat[o:o + n]).reshape(shape)
        o += n
    return out


def _patches(x, k, s):
    """x (B,H,W,C) -> (B,Ho,Wo,k,k,C) view of VALID sliding windows with stride s."""
    B, H, W, C = x.shape
    ho, wo = (H - k) // s + 1, (W - k) // s + 1
    sb, sh, sw, sc = x.strides
    return np.lib.stride_tricks.as_strided(
        x, (B, ho, wo, k, k, C), (sb, sh * s, sw * s, sh, sw, sc), writeable=False)


def conv_fwd(x, w, b, s):
    k = w.shape[0]
    p = _patches(x, k, s)
    B, ho, wo = p.shape[:3]
    cols = p.reshape(B * ho * wo, -1)
    return (cols @ w.reshape(-1, w.shape[3]) + b).reshape(B, ho, wo, -1), cols


def conv_bwd(dy, cols, x_shape, w, s, need_dx=True):
    k, cout = w.shape[0], w.shape[3]
    B, ho, wo = dy.shape[:3]
    d2 = dy.reshape(-1, cout)
    dw = (cols.T @ d2).reshape(w.shape)
    db = d2.sum(0)
    dx = None
    if need_dx:
        dcols = (d2 @ w.reshape(-1, cout).T).reshape(B, ho, wo, k, k, -1)
        dx = np.zeros(x_shape, dtype=dy.dtype)
        for i in range(k):
            for j in range(k):
                dx[:, i:i + s * ho:s, j:j + s * wo:s, :] += dcols[:, :, :, i, j, :]
    return dw, db, dx


def forward(params, s_u8, keep=False):
    """networks.py:181-221,338-363.  s_u8 (B,84,84,4) uint8 -> q (B,A)."""
    dt = params["fc1.w"].dtype
    x = s_u8.astype(dt) / dt.type(255.0)
    cache = {"x0": x}
    for name, _, s, _, _ in LAYERS:
        y, cols = conv_fwd(x, params[name + ".w"], params[name + ".b"], s)
        x = np.maximum(y, 0)
        if keep:
            cache[name + ".cols"], cache[name + ".out"] = cols, x
    h0 = x.reshape(x.shape[0], -1)
    h1 = np.maximum(h0 @ params["fc1.w"] + params["fc1.b"], 0)
    q = h1 @ params["fc2.w"] + params["fc2.b"]      # (1,) bias broadcasts (networks.py:126-132)
    if keep:
        cache["h0"], cache["h1"] = h0, h1
        return q, cache
    return q


def backward(params, cache, dq):
    """Reverse mode through the online net; returns grads in arena order."""
    g = collections.OrderedDict((k, None) for k in params)
    h0, h1 = cache["h0"], cache["h1"]
    g["fc2.w"] = h1.T @ dq
    g["fc2.b"] = dq.sum().reshape(1) if params["fc2.b"].shape == (1,) else dq.sum(0)
    dh1 = (dq @ params["fc2.w"].T) * (h1 > 0)
    cache["dy.fc1"] = dh1
    g["fc1.w"] = h0.T @ dh1
    g["fc1.b"] = dh1.sum(0)
    dx = (dh1 @ params["fc1.w"].T).reshape(cache["conv3.out"].shape)
    shapes_in = {"conv1": cache["x0"].shape, "conv2": cache["conv1.out"].shape,
                 "conv3": cache["conv2.out"].shape}
    for name, _, s, _, _ in reversed(LAYERS):
        dy = dx * (cache[name + ".out"] > 0)
        cache["dy." + name] = dy          # kept for debugging / intermediate parity checks
        dw, db, dx = conv_bwd(dy, cache[name + ".cols"], shapes_in[name],
                              params[name + ".w"], s, need_dx=(name != "conv1"))
        g[name + ".w"], g[name + ".b"] = dw, db
    return g


def conv_dcols(d2, w, x_shape, ho, wo, s):
    """Input gradient of a VALID conv from the flattened cotangent d2 (rows = output positions)."""
    k, cout = w.shape[0], w.shape[3]
    B = x_shape[0]
    dcols = (d2 @ w.reshape(-1, cout).T).reshape(B, ho, wo, k, k, -1)
    dx = np.zeros(x_shape, dtype=d2.dtype)
    for i in range(k):
        for j in range(k):
            dx[:, i:i + s * ho:s, j:j + s * wo:s, :] += dcols[:, :, :, i, j, :]
    return dx


def hvp_backward(params, direction, cache, dq, dq_dot_fn):
    """d/d(eps) of backward(params + eps*direction, ...) at eps = 0: the Hessian-vector product of the loss whose
    cotangent on q is `dq` (forward-over-reverse; ReLU masks are locally constant).

    Tangent forward:  zdot_l = W_l * xdot_{l-1} + Wdot_l * x_{l-1} + bdot_l,  xdot_l = mask_l . zdot_l
    dq_dot_fn(qdot) -> tangent of dq (the loss head: d clip(td/B)/d eps = 1[|td/B| < bound] * (-qdot[a]/B)).
    Tangent backward: ddot_{l-1} = mask . (W_l^T ddot_l + Wdot_l^T d_l),
                      (grad W_l)dot = x_{l-1}^T ddot_l + xdot_{l-1}^T d_l,   (grad b_l)dot = sum ddot_l.
    `cache` must come from forward(params, ..., keep=True) followed by backward() (it holds dy.* per layer)."""
    v = direction
    g = collections.OrderedDict((k, None) for k in params)
    strides = {name: s for name, _, s, _, _ in LAYERS}
    x, xdot = cache["x0"], None
    tcols, shapes_in, tin = {}, {}, {}
    for name, k, s, _, _ in LAYERS:
        w, wd, bd = params[name + ".w"], v[name + ".w"], v[name + ".b"]
        cols = cache[name + ".cols"]
        out = cache[name + ".out"]
        z = cols @ wd.reshape(-1, wd.shape[3]) + bd
        shapes_in[name] = x.shape
        if xdot is not None:
            pd = _patches(xdot, k, s)
            tcols[name] = pd.reshape(cols.shape[0], -1)
            z = z + tcols[name] @ w.reshape(-1, w.shape[3])
        tin[name] = xdot
        xdot = (z.reshape(out.shape)) * (out > 0)
        x = out
    h0, h1 = cache["h0"], cache["h1"]
    h0dot = xdot.reshape(h0.shape)
    h1dot = (h0dot @ params["fc1.w"] + h0 @ v["fc1.w"] + v["fc1.b"]) * (h1 > 0)
    qdot = h1dot @ params["fc2.w"] + h1 @ v["fc2.w"] + v["fc2.b"]
    dqd = dq_dot_fn(qdot)
    g["fc2.w"] = h1dot.T @ dq + h1.T @ dqd
    g["fc2.b"] = dqd.sum().reshape(1) if params["fc2.b"].shape == (1,) else dqd.sum(0)
    dh1 = cache["dy.fc1"]
    dh1d = (dqd @ params["fc2.w"].T + dq @ v["fc2.w"].T) * (h1 > 0)
    g["fc1.w"] = h0dot.T @ dh1 + h0.T @ dh1d
    g["fc1.b"] = dh1d.sum(0)
    dxd = (dh1d @ params["fc1.w"].T + dh1 @ v["fc1.w"].T).reshape(cache["conv3.out"].shape)
    for name, k, s, _, _ in reversed(LAYERS):
        w, wd = params[name + ".w"], v[name + ".w"]
        out = cache[name + ".out"]
        dy = cache["dy." + name]
        dyd = dxd * (out > 0)
        B, ho, wo, cout = dy.shape
        d2, d2d = dy.reshape(-1, cout), dyd.reshape(-1, cout)
        gw = cache[name + ".cols"].T @ d2d
        if name in tcols:
            gw = gw + tcols[name].T @ d2
        g[name + ".w"] = gw.reshape(w.shape)
        g[name + ".b"] = d2d.sum(0)
        if name != "conv1":
            dxd = conv_dcols(d2d, w, shapes_in[name], ho, wo, s) + conv_dcols(d2, wd, shapes_in[name], ho, wo, s)
    return g
