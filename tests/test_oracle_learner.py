"""Self-checks of the NumPy learner oracle (parity unpinned at the jax boundary):
torch float64 autograd, finite differences, and the reference's two known-answer
fragments (networks_test.py:86-103; run_mgsc_test.py:198-218)."""
import numpy as np
import pytest

from dqn_mgsc_zoo_b200 import synthetic
from oracle import learner_oracle as net
from oracle import update_oracle as upd
from tests import torch_ref


def _setup(seed, B, A, shared):
    rng = np.random.default_rng(seed)
    online = net.init_params(rng, A, shared)
    target = net.init_params(rng, A, shared)
    batch = synthetic.random_batch(rng, B, A)
    return rng, online, target, batch


def test_param_count():
    assert sum(int(np.prod(s)) for s in net.param_shapes(6).values()) == 1687206
    assert sum(int(np.prod(s)) for s in net.param_shapes(18).values()) == 1693362
    assert net.param_shapes(6, True)["fc2.b"] == (1,)


def test_shared_bias_known_answer():
    # networks_test.py:86-103: zero weights, bias 1.23 -> every output is 1.23
    rng = np.random.default_rng(0)
    p = net.init_params(rng, 4, shared_bias=True)
    p["fc2.w"][:] = 0
    p["fc2.b"][:] = 1.23
    q = net.forward(p, rng.integers(0, 256, (3, 84, 84, 4), dtype=np.uint8))
    np.testing.assert_allclose(q, 1.23)


def test_softmax_weights_known_answer():
    # run_mgsc_test.py:198-218: softmax-weighted sum of all-ones "gradients" is 1
    rng = np.random.default_rng(1)
    p = upd.softmax(rng.normal(size=17))
    assert abs(p.sum() - 1.0) < 1e-12
    assert abs(p @ np.ones(17) - 1.0) < 1e-12


@pytest.mark.parametrize("double_q,shared,A", [(False, False, 6), (True, True, 18)])
def test_grads_match_torch_autograd(double_q, shared, A):
    _, online, target, batch = _setup(2, 8, A, shared)
    out = upd.loss_and_grads(online, target, batch, double_q)
    loss, q, g = torch_ref.grads(online, target, batch, double_q)
    assert abs(out["loss"] - loss) < 1e-12 * max(1, abs(loss))
    np.testing.assert_allclose(out["q_tm1"], q, rtol=1e-11, atol=1e-12)
    for k in g:
        scale = np.abs(g[k]).max() + 1e-30
        assert np.abs(out["grads"][k] - g[k]).max() / scale < 1e-10, k


def test_importance_weighted_grads_match_torch_autograd():
    # prioritized/agent.py:85-115: loss = mean(w * l2_loss(clip_gradient(td))): the weight reaches the clip
    rng, online, target, batch = _setup(6, 8, 6, True)
    s_tm1, a, r, d, s_t = batch
    r = np.array([40.0, -40.0, 0.3, 0.0, 1.0, -1.0, 0.5, 0.2], dtype=np.float32)      # rows 0/1 clip for w ~ 1
    w = np.array([1.0, 0.2, 0.7, 1.0, 0.05, 0.9, 0.4, 1.0])
    out = upd.loss_and_grads(online, target, (s_tm1, a, r, d, s_t), True, weights=w)
    loss, q, g = torch_ref.grads(online, target, (s_tm1, a, r, d, s_t), True, weights=w)
    assert abs(out["loss"] - loss) < 1e-12 * max(1, abs(loss))
    for k in g:
        scale = np.abs(g[k]).max() + 1e-30
        assert np.abs(out["grads"][k] - g[k]).max() / scale < 1e-10, k
    dq = out["dq"][np.arange(8), a]
    assert dq[0] == -1.0 / 32                      # w = 1: |w td / B| > 1/32 -> clipped
    assert abs(dq[1]) < 1.0 / 32 or dq[1] == 1.0 / 32


def test_clip_is_on_cotangent():
    # with big rewards |td|/B exceeds 1/32 -> dq saturates at +-1/32 (dqn/agent.py:101-103)
    _, online, target, batch = _setup(3, 4, 6, False)
    s_tm1, a, r, d, s_t = batch
    r = np.array([100.0, -100.0, 0.01, 0.0], dtype=np.float32)
    out = upd.loss_and_grads(online, target, (s_tm1, a, r, d, s_t))
    dq = out["dq"][np.arange(4), a]
    assert dq[0] == -1.0 / 32 and dq[1] == 1.0 / 32
    np.testing.assert_allclose(dq[2:], -out["td"][2:] / 4)


def test_finite_difference():
    _, online, target, batch = _setup(4, 2, 6, False)
    s_tm1, a, r, d, s_t = batch
    g = upd.loss_and_grads(online, target, batch, grad_bound=1e9)["grads"]
    rng = np.random.default_rng(0)
    for key in ("conv1.w", "conv2.b", "conv3.w", "fc1.w", "fc2.w", "fc2.b"):
        idx = tuple(rng.integers(0, n) for n in online[key].shape)
        h = 1e-6
        vals = []
        for sign in (+1, -1):
            pert = {k: v.copy() for k, v in online.items()}
            pert[key][idx] += sign * h
            vals.append(upd.loss_and_grads(pert, target, batch, grad_bound=1e9)["loss"])
        fd = (vals[0] - vals[1]) / (2 * h)
        assert abs(fd - g[key][idx]) < 1e-6 * max(1.0, abs(fd)), key


def test_rmsprop_matches_jaxpr_order():
    # Digraph.gv:711-752: mu=(1-d)g+d*mu ; nu=(1-d)g^2+d*nu ; p += -lr*g*rsqrt(nu-mu^2+eps)
    g = np.array([0.5, -2.0]); mu = np.array([0.1, 0.2]); nu = np.array([1.0, 5.0]); p = np.array([1.0, 1.0])
    p2, mu2, nu2 = upd.rmsprop_step(p, g, mu, nu)
    np.testing.assert_allclose(mu2, 0.05 * g + 0.95 * mu)
    np.testing.assert_allclose(nu2, 0.05 * g * g + 0.95 * nu)
    np.testing.assert_allclose(p2, p - 0.00025 * g / np.sqrt(nu2 - mu2 ** 2 + 0.01 / 1024))


def test_adam_first_step_is_lr_sign():
    x, m, v, c = upd.adam_step(np.zeros(3), np.array([1.0, -2.0, 0.5]), np.zeros(3), np.zeros(3), 0)
    np.testing.assert_allclose(x, -0.00025 * np.sign([1.0, -2.0, 0.5]), rtol=1e-6)
    assert c == 1


def test_meta_gradient_matches_autograd():
    rng, online, target, _ = _setup(5, 1, 6, False)
    M = 3
    meta_batch = synthetic.random_batch(rng, M, 6)
    ot = tuple(x[0] for x in synthetic.random_batch(rng, 1, 6))
    P = net.flatten(online).size
    mu = rng.normal(size=P) * 1e-3
    nu = mu ** 2 + rng.random(P) * 1e-4
    logits = rng.normal(size=M)
    out = upd.meta_update(logits, meta_batch, online, target, ot, mu, nu,
                          np.zeros(M), np.zeros(M), 0)
    L, dlogits = torch_ref.meta_grad(logits, meta_batch, online, target, ot, mu, nu)
    assert abs(out["loss"] - L) < 1e-10 * max(1.0, abs(L))
    scale = np.abs(dlogits).max()
    assert scale > 0
    assert np.abs(out["dL_dlogits"] - dlogits).max() / scale < 1e-8


@pytest.mark.parametrize("double_q,shared,A,B", [(False, False, 6, 1), (True, True, 4, 5)])
def test_hvp_matches_torch_double_backward(double_q, shared, A, B):
    rng, online, target, _ = _setup(7, 1, A, shared)
    batch = synthetic.random_batch(rng, B, A)
    batch[2][0] = 40.0 if B > 1 else 0.3         # one clipped row in the batched case: its Gauss-Newton term vanishes
    direction = {k: rng.normal(size=v.shape) * (1.0 / np.sqrt(v.size)) for k, v in online.items()}
    w = rng.random(B) if B > 1 else None
    got = net.unflatten(upd.hvp(online, target, batch, direction, double_q, weights=w), A, shared)
    want = torch_ref.hvp(online, target, batch, direction, double_q, weights=w)
    for k in want:
        scale = np.abs(want[k]).max() + 1e-30
        assert np.abs(got[k] - want[k]).max() / scale < 1e-9, k


def test_second_order_meta_gradient_matches_autograd():
    # dqn_mgsc_batched_reservoir/agent.py:160-198: autograd through g'(theta_exp), mu1, nu1 (no stop_gradient)
    rng, online, target, _ = _setup(8, 1, 6, False)
    M = 3
    meta_batch = synthetic.random_batch(rng, M, 6)
    ot = tuple(x[0] for x in synthetic.random_batch(rng, 1, 6))
    P = net.flatten(online).size
    mu = rng.normal(size=P) * 1e-3
    nu = mu ** 2 + rng.random(P) * 1e-4
    logits = rng.normal(size=M)
    out = upd.meta_update(logits, meta_batch, online, target, ot, mu, nu, np.zeros(M), np.zeros(M), 0,
                          second_order=True)
    L, dlogits = torch_ref.meta_grad(logits, meta_batch, online, target, ot, mu, nu, second_order=True)
    assert abs(out["loss"] - L) < 1e-10 * max(1.0, abs(L))
    scale = np.abs(dlogits).max()
    assert scale > 0
    assert np.abs(out["dL_dlogits"] - dlogits).max() / scale < 1e-7
    first = upd.meta_update(logits, meta_batch, online, target, ot, mu, nu, np.zeros(M), np.zeros(M), 0)
    assert np.abs(first["dL_dlogits"] - out["dL_dlogits"]).max() / scale > 1e-3      # the two objectives differ
