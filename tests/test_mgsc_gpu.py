"""GPU parity of the MGSC path through the C ABI:
  * device logit ring + MGSC replays (dqnb_logits_*) against the golden fixtures made by the
    unmodified reference classes and against the oracle on larger seeded streams
    (indices exact, logits to float32 rounding);
  * dqnb_mgsc_meta_update against oracle.update_oracle.meta_update (float64 restatement of
    dqn_mgsc_batched/agent.py:152-220);
  * the MGSCDqn agent's step ordering (meta-learn before add, learn after).
"""
import numpy as np
import pytest

from dqn_mgsc_zoo_b200 import agent as agent_lib
from dqn_mgsc_zoo_b200 import learner as lrn
from dqn_mgsc_zoo_b200 import networks
from dqn_mgsc_zoo_b200 import parts
from dqn_mgsc_zoo_b200 import replay as rp
from dqn_mgsc_zoo_b200 import replay_circular as rc
from dqn_mgsc_zoo_b200 import replay_jax as rj
from dqn_mgsc_zoo_b200 import synthetic
from oracle import learner_oracle as net
from oracle import replay_oracle as ro
from oracle import update_oracle as upd
from tests import replay_scenarios as sc
from tests.test_oracle_replay import GOLDEN, check_against_golden

pytestmark = pytest.mark.gpu

MGSC = [n for n, s in sc.SCENARIOS.items() if s["kind"].startswith("mgsc")]
PRODUCT = {"mgsc_fifo": rc.MGSCFiFoTransitionReplay, "mgsc_reservoir": rc.MGSCReservoirTransitionReplay}
ORACLE = {"mgsc_fifo": ro.MgscFifoReplay, "mgsc_reservoir": ro.MgscReservoirReplay}


def rel(x, ref):
    x, ref = np.asarray(x, dtype=np.float64), np.asarray(ref, dtype=np.float64)
    return np.abs(x - ref).max() / (np.abs(ref).max() + 1e-30)


@pytest.mark.parametrize("name", MGSC)
def test_mgsc_replay_matches_reference_golden(name):
    spec = sc.SCENARIOS[name]
    got = sc.pairs_mgsc(lambda cap, st, rng: PRODUCT[spec["kind"]](cap, st, rng), **spec["args"])
    # every index stream (weighted draws, meta batches, the caller's next draw) is exact; logits are
    # lse(all) - log(size) chains in float32: the device's expf/logf/summation order give a 1-2 ulp
    # different lse (|lse| ~ 3 here, ulp 2.4e-7), hence the absolute tolerance.
    logits = got.pop("meta_logits")
    check_against_golden(name, got)
    np.testing.assert_allclose(logits, GOLDEN[name + "/meta_logits"], rtol=0, atol=2e-6)


@pytest.mark.parametrize("kind", ["mgsc_fifo", "mgsc_reservoir"])
def test_mgsc_replay_frames_match_oracle(kind):
    """Atari-shaped items: weighted batches bit-exact, logits to float32 rounding, capacity not a
    multiple of the CDF block, several wrap-arounds / reservoir replacements."""
    cap, n_add, batch, meta = 1500, 4000, 32, 10
    o = ORACLE[kind](cap, ro.Transition(None, None, None, None, None), np.random.default_rng(21))
    p = PRODUCT[kind](cap, rp.Transition(None, None, None, None, None), np.random.default_rng(21))
    so = synthetic.transition_stream(np.random.default_rng(4), 6, ro.PairAccumulator(), hw=(84, 84),
                                     short_every=7, max_len=40)
    sp = synthetic.transition_stream(np.random.default_rng(4), 6, rp.TransitionAccumulator(), hw=(84, 84),
                                     short_every=7, max_len=40)
    checks = 0
    for i in range(n_add):
        o.add(next(so))
        p.add(next(sp))
        if i % 97 == 96 and o.size >= batch:
            io, to, lo = o.batch_of_ids_transitions_and_logits(meta)
            ip, tp, lp = p.batch_of_ids_transitions_and_logits(meta)
            assert np.array_equal(io, ip)
            np.testing.assert_allclose(lp, lo, rtol=2e-6, atol=1e-6)
            for x, y in zip(to, tp):
                assert np.array_equal(x, y)
            new = np.asarray(lo, np.float32) + np.float32(0.3) * np.sin(np.arange(meta)).astype(np.float32)
            o.update_priorities(io, new)
            p.update_priorities(ip, new)
            a, b = o.sample(batch), p.sample(batch)
            for x, y, f in zip(a, b, a._fields):
                assert np.array_equal(x, y), (i, f)
            checks += 1
    assert checks >= 30 and p.size == o.size == cap
    np.testing.assert_allclose(p._distribution._logits, o.logits.l if kind == "mgsc_fifo" else o.l,
                               rtol=2e-6, atol=1e-6)
    # the Generator is shared with the caller: both consumed the same number of draws
    assert p._distribution._rng_state.integers(1 << 30) == (o.logits.rng if kind == "mgsc_fifo" else o.rng).integers(1 << 30)


def test_logit_ring_errors_and_state():
    ring = rc.CircularLogitBuffer(5, np.random.default_rng(0))
    with pytest.raises(BufferError, match="empty and cannot be popped"):
        ring.popleft()
    with pytest.raises(BufferError, match="empty and cannot be indexed"):
        ring[np.array([0])]
    for _ in range(5):
        ring.add()
    with pytest.raises(BufferError, match="full and cannot be added"):
        ring.add()
    with pytest.raises(KeyError):
        ring[np.array([5])]
    with pytest.raises(BufferError, match="Cannot sample"):
        rc.CircularLogitBuffer(5, np.random.default_rng(0)).sample(1)
    assert np.allclose(ring._logits, 0.0, atol=1e-6)      # mean-probability logits of an all-equal ring
    ring[np.array([1, 3])] = np.array([2.0, -1.0], np.float32)
    assert np.allclose(ring[np.array([1, 3])], [2.0, -1.0])
    first = ring.popleft()
    assert abs(first) < 1e-6 and ring.size == 4 and np.isneginf(ring._logits[0])
    state = ring.get_state()
    other = rc.CircularLogitBuffer(5, np.random.default_rng(1))
    other.set_state(state)
    assert np.array_equal(other._logits, ring._logits) and other.size == 4 and other._left_head == 1
    # `_logits` is poked directly by the reference's experiments: whole-array assignment works
    ring._logits = np.array([-np.inf, 0.0, 1.0, 2.0, 3.0], np.float32)
    idx = np.concatenate([ring.sample(4) for _ in range(1000)])
    counts = np.bincount(idx, minlength=5)[:4] / 4000.0
    want = np.exp([0.0, 1.0, 2.0, 3.0]) / np.exp([0.0, 1.0, 2.0, 3.0]).sum()
    assert np.abs(counts - want).max() < 0.03
    # replay_jax interface (R15): same method set, key-constructed
    jr = rj.MGSCFiFoTransitionReplay(8, sc.Pair(None, None), rng_key=np.array([0, 7], np.uint32))
    for i in range(12):
        jr.add(sc.Pair(i, -i))
    tr, lg = jr.transitions_and_logits()
    assert np.array_equal(tr.a, np.arange(4, 12)) and lg.shape == (8,)
    with pytest.raises(NotImplementedError):
        jr.slice_of_ids_transitions_and_logits(2)


def _meta_case(A, shared, M, seed):
    rng = np.random.default_rng(seed)
    params = networks.init_params(rng, A, shared)
    target = {k: (v + 0.01 * rng.standard_normal(v.shape)).astype(np.float32) for k, v in params.items()}
    P = net.flatten(params).size
    mu = (1e-3 * rng.standard_normal(P)).astype(np.float32)
    nu = (mu.astype(np.float64) ** 2 + 1e-5 * rng.random(P)).astype(np.float32)
    batch = synthetic.random_batch(rng, M, A)
    one = synthetic.random_batch(rng, 1, A)
    online_tr = (one[0][0], int(one[1][0]), float(one[2][0]), float(one[3][0]), one[4][0])
    logits = rng.normal(0.0, 1.0, M).astype(np.float32)
    return params, target, mu, nu, batch, online_tr, logits


def conditioned_meta_update(meta, logits, batch, params, target, online_tr, mu, nu, adam_m, adam_v, count):
    """oracle.update_oracle.meta_update with every backward pass run through the DEVICE's own forward
    (activations => ReLU masks, td errors) of the batch-of-M and batch-of-1 learners - tests/cond.py.  Same
    formulas, same order (dqn_mgsc_batched/agent.py:160-220); only the discontinuous decisions are the device's."""
    from tests import cond
    A, shared = params["fc2.w"].shape[1], params["fc2.b"].shape == (1,)
    M = len(batch[1])
    size = {110: 512, 111: 3136, 112: 5184, 113: 12800}
    f64 = {k: v.astype(np.float64) for k, v in params.items()}
    p = upd.softmax(np.asarray(logits, np.float64))
    cache = cond.device_cache(lambda tap: meta.read_learner(0, tap, M * size[tap]), batch[0], M)
    td = meta.read_learner(0, lrn.TD, M).astype(np.float64)
    per = np.stack([cond.grads(f64, cond.row_cache(cache, i), td[i:i + 1], batch[1][i:i + 1], 1) for i in range(M)])
    G = p @ per
    theta = net.flatten(f64)
    theta_exp, mu1, nu1 = upd.rmsprop_step(theta, G, mu, nu)
    cache1 = cond.device_cache(lambda tap: meta.read_learner(1, tap, size[tap]), online_tr[0][None], 1)
    td1 = meta.read_learner(1, lrn.TD, 1).astype(np.float64)
    g2 = cond.grads(net.unflatten(theta_exp, A, shared), cache1, td1, [online_tr[1]], 1)
    theta_tgt, _, _ = upd.rmsprop_step(theta_exp, g2, mu1, nu1)
    h = upd.DQN_RMSPROP
    e = 2.0 * (theta_exp - theta_tgt)
    sdev = 1.0 / np.sqrt(nu1 - mu1 * mu1 + h.eps)
    v = e * (-h.lr * sdev * (1.0 - (1.0 - h.decay) * G * (G - mu1) * sdev * sdev))
    dL_dp = per @ v
    new_logits, m2, v2, c2 = upd.adam_step(np.asarray(logits, np.float64), p * (dL_dp - np.dot(p, dL_dp)),
                                           adam_m, adam_v, count)
    return {"logits": new_logits, "adam_m": m2, "adam_v": v2, "adam_count": c2,
            "loss": np.sum((theta_exp - theta_tgt) ** 2), "p": p, "G": G, "v": v, "dL_dp": dL_dp,
            "theta_exp": theta_exp}


# fp32 is compared RAW with the float64 oracle at 2e-5.  tf32 is compared (a) conditioned on the device's own
# forward decisions (conditioned_meta_update) at north_star's rel 1e-3 for every internal quantity (measured:
# G 4e-5, theta_exp 5e-5, v 1.1e-4, dL/dp 4.3e-4, loss 3e-5), and (b) raw at 1e-1: a handful of ReLU units whose
# pre-activation is within tf32 rounding of zero flip, which moves single-sample gradients by ~1e-2, and the
# meta-gradient is an inner product of such gradients.
TF32_META = {"p": 1e-5, "G": 1e-3, "theta_exp": 1e-3, "v": 1e-3, "dL_dp": 1e-3, "loss": 1e-3}


@pytest.mark.parametrize("precision,tol", [("fp32", 2e-5), ("tf32", 1e-1)])
@pytest.mark.parametrize("A,shared,M", [(6, False, 5), (4, True, 9), (6, False, 40)])
def test_meta_update_matches_oracle(precision, tol, A, shared, M):
    params, target, mu, nu, batch, online_tr, logits = _meta_case(A, shared, M, 17 + M)
    L = lrn.Learner(A, 32, shared_bias=shared, double_q=False, precision=precision)
    L.set_params(params, lrn.ONLINE)
    L.set_params(target, lrn.TARGET)
    L.write(lrn.MU, mu)
    L.write(lrn.NU, nu)
    meta = lrn.MetaLearner(L, M)
    f64 = lambda d: {k: v.astype(np.float64) for k, v in d.items()}
    adam_m, adam_v, count = np.zeros(M), np.zeros(M), 0
    cur = logits.copy()
    for it in range(3):     # three chained meta updates: Adam state carries over
        want = upd.meta_update(cur.astype(np.float64), batch, f64(params), f64(target), online_tr,
                               mu.astype(np.float64), nu.astype(np.float64), adam_m, adam_v, count)
        got, loss = meta.meta_update(batch, cur, online_tr)
        errs = {"p": rel(meta.read(lrn.MGSC_P), want["p"]), "G": rel(meta.read(lrn.MGSC_G), want["G"]),
                "theta_exp": rel(meta.read(lrn.MGSC_THETA_EXP) - net.flatten(params),
                                 want["theta_exp"] - net.flatten(f64(params))),
                "v": rel(meta.read(lrn.MGSC_V), want["v"]),
                "dL_dp": rel(meta.read(lrn.MGSC_DL_DP), want["dL_dp"]), "loss": abs(loss - want["loss"]) / want["loss"]}
        print("meta %s A%d M%d it%d:" % (precision, A, M, it), {k: "%.1e" % v for k, v in errs.items()})
        for k, v in errs.items():
            assert v < tol, (it, k, v)
        if precision == "tf32":
            cw = conditioned_meta_update(meta, cur, batch, params, target, online_tr, mu.astype(np.float64),
                                         nu.astype(np.float64), adam_m, adam_v, count)
            cerrs = {"p": rel(meta.read(lrn.MGSC_P), cw["p"]), "G": rel(meta.read(lrn.MGSC_G), cw["G"]),
                     "theta_exp": rel(meta.read(lrn.MGSC_THETA_EXP) - net.flatten(params),
                                      cw["theta_exp"] - net.flatten(f64(params))),
                     "v": rel(meta.read(lrn.MGSC_V), cw["v"]), "dL_dp": rel(meta.read(lrn.MGSC_DL_DP), cw["dL_dp"]),
                     "loss": abs(loss - cw["loss"]) / cw["loss"]}
            print("meta tf32 conditioned A%d M%d it%d:" % (A, M, it), {k: "%.1e" % v for k, v in cerrs.items()})
            for k, v in cerrs.items():
                assert v < TF32_META[k], (it, k, v)
            assert np.abs(got - cw["logits"]).max() < 5.1e-4      # one Adam step is 2.5e-4 whatever |g| is: a sign flip of a ~0 entry
        # logits are float32 of magnitude ~1 and one Adam step is 2.5e-4: compare the values
        assert np.abs(got - want["logits"]).max() < (2e-7 if precision == "fp32" else 5e-4)
        adam_m, adam_v, count = want["adam_m"], want["adam_v"], want["adam_count"]
        cur = got
    st = meta.get_state()
    assert st["count"] == 3
    np.testing.assert_allclose(st["mu"], adam_m, rtol=100 * tol, atol=1e-12 if precision == "fp32" else 1e-6)
    # virtual steps only: the learner's params and RMSProp state did not move
    assert np.array_equal(L.read(lrn.ONLINE), net.flatten(params))
    assert np.array_equal(L.read(lrn.MU), mu) and np.array_equal(L.read(lrn.NU), nu)
    # the replay-slot entry point gathers the same batch on the device
    store = rp.TransitionReplay(64, rp.Transition(None, None, None, None, None), np.random.RandomState(0))
    for i in range(M):
        store.add(rp.Transition(batch[0][i], int(batch[1][i]), float(batch[2][i]), float(batch[3][i]), batch[4][i]))
    meta2 = lrn.MetaLearner(L, M)
    a, la = meta2.meta_update_from_replay(store.frame_store, np.arange(M), logits, online_tr)
    meta3 = lrn.MetaLearner(L, M)
    b, lb = meta3.meta_update(batch, logits, online_tr)
    assert np.array_equal(a, b) and la == lb


def test_mgsc_agent_schedule_and_state():
    A, B, M, CAP = 6, 8, 5, 64
    replay = rc.MGSCFiFoTransitionReplay(CAP, rp.Transition(None, None, None, None, None), np.random.default_rng(3))
    ag = agent_lib.MGSCDqn(
        preprocessor=synthetic.StackingPreprocessor(num_action_repeats=2),
        sample_network_input=np.zeros((84, 84, 4), np.uint8), network=networks.dqn_atari_network(A),
        optimizer=networks.rmsprop(learning_rate=0.00025, decay=0.95, eps=0.01 / 32 ** 2, centered=True),
        transition_accumulator=rp.TransitionAccumulator(), replay=replay, batch_size=B,
        exploration_epsilon=lambda t: 1.0, min_replay_capacity_fraction=0.25, learn_period=4,
        target_network_update_period=24, grad_error_bound=1.0 / 32, rng_key=5,
        meta_optimizer=networks.adam(0.00025), meta_batch_size=M)
    events = []
    real_meta, real_add, real_learn = ag._meta_prioritization_learn, replay.add, ag._learn
    ag._meta_prioritization_learn = lambda tr: (events.append(("meta", ag._frame_t, replay.size)), real_meta(tr))[1]
    replay.add = lambda item: (events.append(("add", ag._frame_t, replay.size)), real_add(item))[1]
    ag._learn = lambda: (events.append(("learn", ag._frame_t, replay.size)), real_learn())[1]
    env = synthetic.SyntheticEnvironment(np.random.default_rng(11), A, min_len=3, max_len=40)
    frames = 0
    for _, timestep, _, _ in parts.run_loop(ag, env, max_steps_per_episode=30, yield_before_reset=True):
        if timestep is None:
            continue
        frames += 1
        if frames >= 200:
            break
    metas = [e for e in events if e[0] == "meta"]
    learns = [e for e in events if e[0] == "learn"]
    assert len(metas) >= 10 and len(learns) >= len(metas)
    for kind, t, size in metas:
        assert t % 4 == 0 and size >= max(M, 0.25 * CAP)
        same_frame = [e[0] for e in events if e[1] == t]
        # dqn_mgsc_batched/agent.py:252-278: meta-learn before the add of that frame, learn after
        assert same_frame.index("meta") < (same_frame.index("add") if "add" in same_frame else 99)
        assert "learn" in same_frame and same_frame.index("meta") < same_frame.index("learn")
    assert np.isfinite(ag.statistics["meta_loss"]) and np.isfinite(ag.statistics["state_value"])
    logits = replay._distribution._logits
    assert np.isfinite(logits[:replay.size]).all() and np.ptp(logits[np.isfinite(logits)]) > 0
    state = ag.get_state()
    assert set(state) == {"rng_key", "frame_t", "opt_state", "online_params", "target_params", "replay",
                          "meta_opt_state"}
    assert state["meta_opt_state"][0].count == len(metas) and state["meta_opt_state"][0].mu.shape == (M,)


def test_logit_ring_at_one_million_entries_matches_reference_semantics():
    """The benchmarked scale: a 1 M-entry ring, 64 weighted draws and the add / popleft logit
    chain against the oracle (= Generator.choice(p=softmax) and lse(all) - log(size), replay_circular.py:166-217).
    The ring is loaded through whole-array state (filling it add by add costs the oracle 2 ms per add); then the
    running logsumexp is exercised across more than one exact re-anchoring (kExactEvery = 64 mutations)."""
    cap, live = 1_000_000, 999_000
    rng = np.random.default_rng(123)
    logits = np.full(cap, -np.inf, dtype=np.float32)
    head = 4321
    pos = (head + np.arange(live)) % cap
    logits[pos] = rng.normal(0.0, 1.5, live).astype(np.float32)
    o = ro.RingLogits(cap, np.random.default_rng(9))
    o.l[:] = logits
    o.size, o.head, o.tail = live, head, (head + live) % cap
    p = rc.CircularLogitBuffer(cap, np.random.default_rng(9))
    p.set_state({'capacity': cap, 'logits': logits.copy(), 'size': live, 'left_head': head,
                 'right_head': (head + live) % cap, 'rng_state': np.random.default_rng(9)})
    assert np.array_equal(p.sample(64), o.sample(64))
    for step in range(150):                       # crosses two exact passes of the running logsumexp
        if step % 3 == 2:
            a, b = p.popleft(), o.popleft()
            assert abs(float(a) - float(b)) == 0.0
        else:
            p.add()
            o.push()
        if step % 10 == 9:
            k = np.array([p.size - 1, p.size - 2, 0])
            np.testing.assert_allclose(p[k], o.get(k), rtol=0, atol=2e-6)
        if step % 50 == 49:
            idx = np.array([5, 17, 99, 4000])
            vals = np.array([3.0, -2.0, 9.5, 0.25], np.float32)      # 9.5 becomes the new maximum
            p[idx] = vals
            o.set(idx, vals)
            assert np.array_equal(p.sample(64), o.sample(64)), step
    np.testing.assert_allclose(p._logits, o.l, rtol=0, atol=2e-6)
    assert np.array_equal(p.sample(64), o.sample(64))
    assert p._rng_state.integers(1 << 30) == o.rng.integers(1 << 30)


@pytest.mark.parametrize("B,A,shared,double_q,weighted", [(1, 6, False, False, 0), (5, 4, True, True, 2),
                                                          (32, 6, False, False, 0)])
def test_hessian_vector_product_matches_oracle(B, A, shared, double_q, weighted):
    """dqnb_learner_hvp (forward-over-reverse through the layer GEMMs) against the oracle R-op, which
    tests/test_oracle_learner.py pins to torch double backward.  fp32 learner; batch 1 is the MGSC use."""
    rng = np.random.default_rng(31 + B)
    params = networks.init_params(rng, A, shared)
    target = {k: (v + 0.01 * rng.standard_normal(v.shape)).astype(np.float32) for k, v in params.items()}
    L = lrn.Learner(A, B, shared_bias=shared, double_q=double_q, precision="fp32", defer_apply=True,
                    row_weighted=weighted)
    L.set_params(params, lrn.ONLINE)
    L.set_params(target, lrn.TARGET)
    batch = synthetic.random_batch(rng, B, A)
    if B > 1:
        batch[2][0] = 40.0                       # a clipped row: its Gauss-Newton term must vanish
    w = rng.random(B).astype(np.float32) if weighted else None
    if weighted:
        L.set_row_weights(w)
    L.update_host(*batch)
    f64 = lambda d: {k: v.astype(np.float64) for k, v in d.items()}
    P = L.num_params
    for trial in range(2):                       # two directions: the scratch of the first call is reused
        v = (rng.standard_normal(P) / np.sqrt(P)).astype(np.float32)
        got = L.hvp(v)
        want = upd.hvp(f64(params), f64(target), batch, v.astype(np.float64), double_q, weights=w)
        got_t, want_t = lrn.unflatten_params(got, A, shared), net.unflatten(want, A, shared)
        errs = {k: rel(got_t[k], want_t[k]) for k in want_t}
        print("hvp B%d trial %d:" % (B, trial), {k: "%.1e" % e for k, e in errs.items()})
        for k, e in errs.items():
            assert e < 5e-5, (k, e)
        assert rel(got, want) < 2e-5
    # the update's own outputs are untouched by the product
    g = upd.loss_and_grads(f64(params), f64(target), batch, double_q,
                           weights=None if w is None else w.astype(np.float64))["grads"]
    assert rel(L.read(lrn.GRAD), net.flatten(g)) < 1e-5


@pytest.mark.parametrize("A,shared,M", [(6, False, 5), (4, True, 9)])
def test_second_order_meta_update_matches_oracle(A, shared, M):
    """The dqn_mgsc_batched_reservoir objective (no stop_gradient, agent.py:160-198) against the oracle's
    second_order=True path (pinned to torch autograd through the whole meta loss in tests/test_oracle_learner.py)."""
    params, target, mu, nu, batch, online_tr, logits = _meta_case(A, shared, M, 41 + M)
    L = lrn.Learner(A, 32, shared_bias=shared, double_q=False, precision="fp32")
    L.set_params(params, lrn.ONLINE)
    L.set_params(target, lrn.TARGET)
    L.write(lrn.MU, mu)
    L.write(lrn.NU, nu)
    meta = lrn.MetaLearner(L, M, second_order=True)
    f64 = lambda d: {k: v.astype(np.float64) for k, v in d.items()}
    adam_m, adam_v, count = np.zeros(M), np.zeros(M), 0
    cur = logits.copy()
    for it in range(2):
        want = upd.meta_update(cur.astype(np.float64), batch, f64(params), f64(target), online_tr,
                               mu.astype(np.float64), nu.astype(np.float64), adam_m, adam_v, count, second_order=True)
        got, loss = meta.meta_update(batch, cur, online_tr)
        errs = {"G": rel(meta.read(lrn.MGSC_G), want["G"]), "v": rel(meta.read(lrn.MGSC_V), want["v"]),
                "dL_dp": rel(meta.read(lrn.MGSC_DL_DP), want["dL_dp"]), "loss": abs(loss - want["loss"]) / want["loss"]}
        print("meta second-order A%d M%d it%d:" % (A, M, it), {k: "%.1e" % v for k, v in errs.items()})
        for k, v in errs.items():
            assert v < 1e-4, (it, k, v)
        assert np.abs(got - want["logits"]).max() < 5e-7
        adam_m, adam_v, count = want["adam_m"], want["adam_v"], want["adam_count"]
        cur = got
    first = upd.meta_update(logits.astype(np.float64), batch, f64(params), f64(target), online_tr,
                            mu.astype(np.float64), nu.astype(np.float64), np.zeros(M), np.zeros(M), 0)
    second = upd.meta_update(logits.astype(np.float64), batch, f64(params), f64(target), online_tr,
                             mu.astype(np.float64), nu.astype(np.float64), np.zeros(M), np.zeros(M), 0, second_order=True)
    assert rel(first["v"], second["v"]) > 1e-2          # the two objectives really differ
    assert np.array_equal(L.read(lrn.ONLINE), net.flatten(params))      # virtual steps only


def test_mgsc_reservoir_agent_runs_the_second_order_objective():
    """dqn_mgsc_batched_reservoir: MGSCReservoirDqn + MGSCReservoirTransitionReplay (run_atari.py:211), a short run."""
    A, B, M, CAP = 6, 8, 5, 48
    replay = rc.MGSCReservoirTransitionReplay(CAP, rp.Transition(None, None, None, None, None), np.random.default_rng(3))
    ag = agent_lib.MGSCReservoirDqn(
        preprocessor=synthetic.StackingPreprocessor(num_action_repeats=2),
        sample_network_input=np.zeros((84, 84, 4), np.uint8), network=networks.dqn_atari_network(A),
        optimizer=networks.rmsprop(learning_rate=0.00025, decay=0.95, eps=0.01 / 32 ** 2, centered=True),
        transition_accumulator=rp.TransitionAccumulator(), replay=replay, batch_size=B,
        exploration_epsilon=lambda t: 1.0, min_replay_capacity_fraction=0.25, learn_period=4,
        target_network_update_period=24, grad_error_bound=1.0 / 32, rng_key=5,
        meta_optimizer=networks.adam(0.00025), meta_batch_size=M, precision="fp32")
    assert ag.meta_variant == 'second_order' and ag.meta_learner.second_order
    assert agent_lib.MGSCDqn._SECOND_ORDER is False
    env = synthetic.SyntheticEnvironment(np.random.default_rng(11), A, min_len=3, max_len=40)
    frames = 0
    for _, timestep, _, _ in parts.run_loop(ag, env, max_steps_per_episode=30, yield_before_reset=True):
        if timestep is None:
            continue
        frames += 1
        if frames >= 200:
            break
    assert replay.size == CAP and replay._t > CAP                 # reservoir replacements happened
    assert np.isfinite(ag.statistics["meta_loss"]) and ag.statistics["meta_loss"] > 0
    logits = replay._distribution._logits
    assert np.isfinite(logits).all() and np.ptp(logits) > 0
    assert ag.get_state()["meta_opt_state"][0].count >= 10
