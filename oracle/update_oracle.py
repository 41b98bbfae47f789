"""TEST INFRASTRUCTURE ONLY — TD loss, optimisers, `update` and MGSC `meta_update`.

PARITY UNPINNED (see oracle/learner_oracle.py header): restates rlax 0.1.5
(q_learning, double_q_learning, clip_gradient, l2_loss) and optax 0.1.2
(scale_by_stddev -> rmsprop centered; scale_by_adam) as used at
  dqn/agent.py:85-117, double_q/agent.py:85-111, dqn/run_atari.py:208-213,
  dqn_mgsc_batched/agent.py:104-220, dqn_mgsc_batched/run_atari.py:241-243,
with op order cross-checked against Digraph.gv:279-340 (TD) and :711-752
(RMSProp: (1-d)*g + d*mu ; eps inside the rsqrt).
"""
import collections

import numpy as np

from oracle import learner_oracle as net

RmsProp = collections.namedtuple("RmsProp", "lr decay eps")
DQN_RMSPROP = RmsProp(lr=0.00025, decay=0.95, eps=0.01 / 32 ** 2)   # dqn/run_atari.py:208-213
Adam = collections.namedtuple("Adam", "lr b1 b2 eps")
META_ADAM = Adam(lr=0.00025, b1=0.9, b2=0.999, eps=1e-8)             # optax.adam defaults


def td_and_dq(q_tm1, a, r, disc, q_target_t, q_sel_t=None, grad_bound=1.0 / 32, mean_over=None, weights=None):
    """td errors, loss and d loss / d q_tm1.

    q_learning: y = r + disc*max_a q_target_t;  double: y = r + disc*q_target_t[argmax q_sel_t]
    td = y - q_tm1[a]; loss = mean(0.5 td^2); clip_gradient clips the *cotangent*
    td/B to +-grad_bound (dqn/agent.py:94-106).  mean_over overrides B (MGSC batch-of-1).
    weights: importance-sampling weights of prioritized replay (prioritized/agent.py:85-115):
    loss = mean(w * 0.5 td^2), so the cotangent reaching clip_gradient is w*td/B.
    """
    B = q_tm1.shape[0]
    rows = np.arange(B)
    sel = np.argmax(q_target_t if q_sel_t is None else q_sel_t, axis=1)
    y = r + disc * q_target_t[rows, sel]
    td = y - q_tm1[rows, a]
    denom = B if mean_over is None else mean_over
    w = 1.0 if weights is None else np.asarray(weights, dtype=q_tm1.dtype)
    cot = np.clip(td * w / denom, -grad_bound, grad_bound)
    dq = np.zeros_like(q_tm1)
    dq[rows, a] = -cot
    loss = np.sum(0.5 * td * td * w) / denom
    return td, loss, dq


def rmsprop_step(p, g, mu, nu, h=DQN_RMSPROP):
    """optax 0.1.2 scale_by_stddev + scale(-lr) + apply_updates, on flat vectors."""
    mu2 = (1 - h.decay) * g + h.decay * mu
    nu2 = (1 - h.decay) * (g * g) + h.decay * nu
    upd = g * (1.0 / np.sqrt(nu2 - mu2 * mu2 + h.eps))
    return p + (-h.lr) * upd, mu2, nu2


def adam_step(x, g, m, v, count, h=META_ADAM):
    """optax 0.1.2 scale_by_adam (bias-corrected, eps outside sqrt) + scale(-lr)."""
    m2 = (1 - h.b1) * g + h.b1 * m
    v2 = (1 - h.b2) * (g * g) + h.b2 * v
    c = count + 1
    mh = m2 / (1 - h.b1 ** c)
    vh = v2 / (1 - h.b2 ** c)
    return x + (-h.lr) * (mh / (np.sqrt(vh) + h.eps)), m2, v2, c


def loss_and_grads(online, target, batch, double_q=False, grad_bound=1.0 / 32, mean_over=None, weights=None):
    """jax.grad(loss_fn) (dqn/agent.py:112-114).  batch = (s_tm1,a,r,disc,s_t)."""
    s_tm1, a, r, disc, s_t = batch
    dt = online["fc1.w"].dtype
    r, disc = np.asarray(r, dtype=dt), np.asarray(disc, dtype=dt)
    q_tm1, cache = net.forward(online, s_tm1, keep=True)
    q_tgt = net.forward(target, s_t)
    q_sel = net.forward(online, s_t) if double_q else None
    td, loss, dq = td_and_dq(q_tm1, np.asarray(a), r, disc, q_tgt, q_sel, grad_bound, mean_over, weights)
    grads = net.backward(online, cache, dq)
    return {"q_tm1": q_tm1, "q_target_t": q_tgt, "q_sel_t": q_sel, "td": td, "loss": loss,
            "dq": dq, "grads": grads, "cache": cache}


def update(online, target, mu, nu, batch, double_q=False, grad_bound=1.0 / 32, h=DQN_RMSPROP, weights=None):
    """`update` of dqn/agent.py:109-117 on flat opt-state vectors mu, nu."""
    A = online["fc2.w"].shape[1]
    shared = online["fc2.b"].shape == (1,)
    out = loss_and_grads(online, target, batch, double_q, grad_bound, weights=weights)
    g = net.flatten(out["grads"])
    p2, mu2, nu2 = rmsprop_step(net.flatten(online), g, mu, nu, h)
    out.update(params=net.unflatten(p2, A, shared), mu=mu2, nu=nu2, flat_grad=g)
    return out


def hvp(online, target, batch, direction, double_q=False, grad_bound=1.0 / 32, mean_over=None, weights=None):
    """H . direction, H = Jacobian of jax.grad(loss_fn) w.r.t. the online params (what differentiating THROUGH
    jax.grad(loss_fn)(expected_online_params, ...) needs, dqn_mgsc_batched_reservoir/agent.py:183-185).
    rlax.clip_gradient's backward `clip(g)` is itself differentiated: 1 inside the bounds, 0 outside; the target
    (and a double-Q selector) carry no gradient.  Returns the flat vector."""
    out = loss_and_grads(online, target, batch, double_q, grad_bound, mean_over, weights)
    s_tm1, a, r, disc, s_t = batch
    a = np.asarray(a)
    B = len(a)
    rows = np.arange(B)
    denom = B if mean_over is None else mean_over
    w = np.ones(B) if weights is None else np.asarray(weights, dtype=np.float64)
    inside = np.abs(out["td"] * w / denom) < grad_bound

    def dq_dot(qdot):
        d = np.zeros_like(qdot)
        d[rows, a] = np.where(inside, qdot[rows, a] * w / denom, 0.0)     # -(d clip(td w / B)) with td_dot = -qdot[a]
        return d
    A = online["fc2.w"].shape[1]
    shared = online["fc2.b"].shape == (1,)
    v = direction if isinstance(direction, dict) else net.unflatten(np.asarray(direction), A, shared)
    return net.flatten(net.hvp_backward(online, v, out["cache"], out["dq"], dq_dot))


def softmax(logits):
    c = logits.max()
    return np.exp(logits - (c + np.log(np.sum(np.exp(logits - c)))))


def meta_update(logits, meta_batch, online, target, online_transition, mu, nu, adam_m, adam_v,
                adam_count, grad_bound=1.0 / 32, h=DQN_RMSPROP, ha=META_ADAM, second_order=False):
    """MGSC meta_update (dqn_mgsc_batched/agent.py:160-220), stop-gradient variant; second_order=True is the
    dqn_mgsc_batched_reservoir/agent.py:160-198 objective (no stop_gradient on target_online_params): the loss
    ||theta_exp - theta_tgt||^2 = ||u'||^2 with u' = rmsprop(g'(theta_exp), (mu1, nu1)) differentiated through g'
    (a Hessian-vector product), mu1 and nu1.

    Materialises the M per-sample gradients exactly like the reference's vmap
    (fine at oracle sizes).  Returns new logits, new Adam state and internals.
    """
    A = online["fc2.w"].shape[1]
    shared = online["fc2.b"].shape == (1,)
    s_tm1, a, r, disc, s_t = meta_batch
    M = len(a)
    p = softmax(np.asarray(logits, dtype=online["fc1.w"].dtype))
    per = []
    for i in range(M):   # weighted_grads: batch-of-1 loss => clip acts on td itself (:152-158)
        one = (s_tm1[i:i + 1], a[i:i + 1], r[i:i + 1], disc[i:i + 1], s_t[i:i + 1])
        per.append(net.flatten(loss_and_grads(online, target, one, False, grad_bound)["grads"]))
    per = np.stack(per)                                   # (M, P)
    G = p @ per
    theta = net.flatten(online)
    theta_exp, mu1, nu1 = rmsprop_step(theta, G, mu, nu, h)              # :178-179
    o = online_transition
    one = tuple(np.asarray(x)[None, ...] for x in o)
    g2 = net.flatten(loss_and_grads(net.unflatten(theta_exp, A, shared), online, one,
                                    False, grad_bound)["grads"])          # :183-185 target=online
    theta_tgt, _, _ = rmsprop_step(theta_exp, g2, mu1, nu1, h)            # :189-191 (stop-grad)
    L = np.sum((theta_exp - theta_tgt) ** 2)                              # :195
    # dL/dG through theta_exp = theta - lr*G*rsqrt(nu1 - mu1^2 + eps), mu1, nu1 functions of G.
    s = 1.0 / np.sqrt(nu1 - mu1 * mu1 + h.eps)
    du_dG = -h.lr * s * (1.0 - (1.0 - h.decay) * G * (G - mu1) * s * s)
    if not second_order:
        e = 2.0 * (theta_exp - theta_tgt)
        v = e * du_dG
    else:
        # L = ||u'||^2, u' = -lr g' s2, s2 = rsqrt(nu2 - mu2^2 + eps), mu2 = (1-d) g' + d mu1, nu2 = (1-d) g'^2 + d nu1
        d = h.decay
        mu2 = (1 - d) * g2 + d * mu1
        nu2 = (1 - d) * g2 * g2 + d * nu1
        s2 = 1.0 / np.sqrt(nu2 - mu2 * mu2 + h.eps)
        e = 2.0 * (theta_tgt - theta_exp)                              # 2 u'
        du_dg = -h.lr * (s2 - (1 - d) * g2 * (g2 - mu2) * s2 ** 3)      # direct + through mu2, nu2
        du_dmu1 = -h.lr * g2 * d * mu2 * s2 ** 3
        du_dnu1 = 0.5 * h.lr * g2 * d * s2 ** 3
        t = hvp(net.unflatten(theta_exp, A, shared), online, one, e * du_dg, False, grad_bound)   # cotangent on theta_exp
        v = t * du_dG + e * du_dmu1 * (1 - d) + e * du_dnu1 * 2 * (1 - d) * G
    dL_dp = per @ v
    dL_dlogits = p * (dL_dp - np.dot(p, dL_dp))
    new_logits, m2, v2, c2 = adam_step(np.asarray(logits, dtype=p.dtype), dL_dlogits,
                                       adam_m, adam_v, adam_count, ha)
    return {"logits": new_logits, "adam_m": m2, "adam_v": v2, "adam_count": c2, "loss": L,
            "p": p, "G": G, "v": v, "dL_dp": dL_dp, "dL_dlogits": dL_dlogits,
            "theta_exp": theta_exp, "theta_tgt": theta_tgt}
