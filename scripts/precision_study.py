"""Emulates reduced-precision GEMM operands (bf16 / tf32, fp32+ accumulate) inside the NumPy
oracle to predict the parity error of a tensor-core learner against the float64 oracle.
Run on CPU:  python scripts/precision_study.py"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from oracle import learner_oracle as net
from oracle import update_oracle as upd
from dqn_mgsc_zoo_b200 import synthetic


def r_bf16(x):
    return torch.from_numpy(np.asarray(x, np.float32)).bfloat16().float().numpy().astype(np.float64)


def r_tf32(x):  # round to nearest, 10 explicit mantissa bits
    x = np.asarray(x, np.float32)
    u = x.view(np.uint32).astype(np.uint64)
    u = (u + 0x1000) & 0xFFFFE000
    return u.astype(np.uint32).view(np.float32).astype(np.float64)


def r_bf16x2(x):  # hi + lo split: ~16 mantissa bits
    hi = r_bf16(x)
    return hi + r_bf16(np.asarray(x, np.float64) - hi)


def run(rf, online, target, batch, double_q, scale_in_epilogue=True):
    s_tm1, a, r, disc, s_t = batch

    def fwd(params, s, keep=False):
        x = s.astype(np.float64)            # raw 0..255: exact in bf16 / tf32
        cache = {}
        first = True
        for name, _, st, _, _ in net.LAYERS:
            w, b = params[name + ".w"], params[name + ".b"]
            p = net._patches(rf(x), w.shape[0], st)
            B, ho, wo = p.shape[:3]
            cols = p.reshape(B * ho * wo, -1)
            y = cols @ rf(w).reshape(-1, w.shape[3])
            if first:
                y = y / 255.0
                first = False
            y = (y + b).reshape(B, ho, wo, -1)
            cache[name + ".in"] = x
            x = np.maximum(y, 0)
            cache[name + ".out"] = x
        h0 = x.reshape(x.shape[0], -1)
        h1 = np.maximum(rf(h0) @ rf(params["fc1.w"]) + params["fc1.b"], 0)
        q = h1 @ params["fc2.w"] + params["fc2.b"]      # fc2 stays fp32 SIMT
        cache["h0"], cache["h1"] = h0, h1
        return (q, cache) if keep else q

    q_tm1, c = fwd(online, s_tm1, True)
    q_tgt = fwd(target, s_t)
    q_sel = fwd(online, s_t) if double_q else None
    td, loss, dq = upd.td_and_dq(q_tm1, np.asarray(a), np.asarray(r, np.float64), np.asarray(disc, np.float64), q_tgt, q_sel)
    g = {}
    h0, h1 = c["h0"], c["h1"]
    g["fc2.w"] = h1.T @ dq
    g["fc2.b"] = dq.sum().reshape(1) if online["fc2.b"].shape == (1,) else dq.sum(0)
    dh1 = (dq @ online["fc2.w"].T) * (h1 > 0)
    g["fc1.w"] = rf(h0).T @ rf(dh1)
    g["fc1.b"] = dh1.sum(0)
    dx = (rf(dh1) @ rf(online["fc1.w"]).T).reshape(c["conv3.out"].shape)
    for name, _, st, _, _ in reversed(net.LAYERS):
        w = online[name + ".w"]
        dy = dx * (c[name + ".out"] > 0)
        xin = c[name + ".in"]
        p = net._patches(rf(xin), w.shape[0], st)
        B, ho, wo = p.shape[:3]
        cols = p.reshape(B * ho * wo, -1)
        d2 = rf(dy).reshape(-1, w.shape[3])
        dw = (cols.T @ d2).reshape(w.shape)
        if name == "conv1":
            dw = dw / 255.0
        g[name + ".w"], g[name + ".b"] = dw, dy.reshape(-1, w.shape[3]).sum(0)
        if name != "conv1":
            k = w.shape[0]
            dcols = (d2 @ rf(w).reshape(-1, w.shape[3]).T).reshape(B, ho, wo, k, k, -1)
            dx = np.zeros(xin.shape)
            for i in range(k):
                for j in range(k):
                    dx[:, i:i + st * ho:st, j:j + st * wo:st, :] += dcols[:, :, :, i, j, :]
    return q_tm1, q_tgt, td, loss, g


def rel(x, ref):
    return np.abs(np.asarray(x) - ref).max() / (np.abs(ref).max() + 1e-30)


def main():
    A, B = 6, 32
    for seed in (0, 1):
        rng = np.random.default_rng(seed)
        online = net.init_params(rng, A, True)
        target = net.init_params(rng, A, True)
        batch = synthetic.random_batch(rng, B, A)
        want = upd.loss_and_grads(online, target, batch, True)
        for nm, rf in (("bf16", r_bf16), ("tf32", r_tf32), ("bf16x2", r_bf16x2), ("exact", lambda x: np.asarray(x, np.float64))):
            q, qt, td, loss, g = run(rf, online, target, batch, True)
            errs = {k: rel(g[k], want["grads"][k]) for k in g}
            print(seed, nm, "q %.2e qt %.2e td %.2e loss %.2e | grads max %.2e" % (
                rel(q, want["q_tm1"]), rel(qt, want["q_target_t"]), rel(td, want["td"]),
                abs(loss - want["loss"]) / abs(want["loss"]), max(errs.values())),
                {k: "%.1e" % v for k, v in errs.items()})


if __name__ == "__main__":
    main()
