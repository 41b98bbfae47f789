"""Pins oracle/replay_oracle.py: (1) against the committed golden fixtures made by
the unmodified reference (tests/golden/make_golden.py); (2) when the reference
tree is present (build container), against the reference classes directly on
fresh seeds; (3) the reference's own statistical reservoir test
(dqn_reservoir/run_reservoir_test.py:26-47) and API invariants
(replay_test.py:99-174)."""
import os

import numpy as np
import pytest

from dqn_mgsc_zoo_b200 import synthetic
from oracle import ref_shim
from oracle import replay_oracle as ro
from tests import replay_scenarios as sc

GOLDEN = np.load(os.path.join(os.path.dirname(__file__), "golden", "replay_golden.npz"))


def oracle_factory(kind):
    if kind in ("per", "gen_per"):
        return lambda capacity, structure, alpha, beta, usp, normalize, rng: ro.PrioritizedReplay(
            capacity, structure, alpha, beta, usp, normalize, rng)
    cls = {"fifo": ro.FifoReplay, "reservoir": ro.ReservoirReplay,
           "gen_fifo": ro.FifoReplay, "gen_reservoir": ro.ReservoirReplay,
           "mgsc_fifo": ro.MgscFifoReplay, "mgsc_reservoir": ro.MgscReservoirReplay}[kind]
    return lambda capacity, structure, rng: cls(capacity, structure, rng)


def run(name, factory, accumulator, structure):
    spec = sc.SCENARIOS[name]
    fn = getattr(sc, spec["fn"])
    if spec["fn"] == "frames_uniform":
        return fn(factory(spec["kind"]), accumulator, structure, **spec["args"])
    return fn(factory(spec["kind"]), **spec["args"])


def check_against_golden(name, got):
    for key, val in got.items():
        want = GOLDEN[name + "/" + key]
        val = np.asarray(val)
        if val.dtype.kind == "f" and "logits" in key:
            np.testing.assert_allclose(val, want, rtol=1e-6, atol=1e-7, err_msg=name + "/" + key)
        else:
            assert np.array_equal(val, want), name + "/" + key


@pytest.mark.parametrize("name", list(sc.SCENARIOS))
def test_oracle_matches_golden(name):
    got = run(name, oracle_factory, ro.PairAccumulator(), ro.Transition(None, None, None, None, None))
    check_against_golden(name, got)


@pytest.mark.skipif(not ref_shim.available(), reason="reference tree not on this machine")
@pytest.mark.parametrize("kind", ["fifo", "reservoir"])
def test_oracle_matches_reference_fresh_seed(kind):
    from tests.golden.make_golden import reference_factory
    args = dict(seed=1234, capacity=29, n_add=400, sample_every=2, batch=7)
    a = sc.pairs_uniform(reference_factory(kind), **args)
    b = sc.pairs_uniform(oracle_factory(kind), **args)
    for k in a:
        assert np.array_equal(a[k], b[k]), k


@pytest.mark.skipif(not ref_shim.available(), reason="reference tree not on this machine")
@pytest.mark.parametrize("kind", ["per", "gen_per"])
def test_oracle_prioritized_matches_reference_fresh_seed(kind):
    from tests.golden.make_golden import reference_factory
    args = dict(seed=4321, capacity=21, n_add=200, every=2, batch=5, alpha=0.7, usp=0.05, generator=kind == "gen_per",
                zero_until=8)
    a = sc.pairs_prioritized(reference_factory(kind), **args)
    b = sc.pairs_prioritized(oracle_factory(kind), **args)
    for k in a:
        assert np.array_equal(a[k], b[k]), k


def test_prefix_tree_known_answers():
    # SumTreeTest (replay_test.py:820-1045): queries return the smallest index whose running sum exceeds the target
    t = ro.PrefixTree(5)
    t.assign(range(5), [1.0, 0.0, 2.5, 0.5, 4.0])
    assert t.total == 8.0 and t.width == 8
    assert [t.locate(x) for x in (0.0, 0.99, 1.0, 3.49, 3.5, 3.99, 4.0, 7.99)] == [0, 0, 2, 2, 3, 3, 4, 4]
    with pytest.raises(ValueError):
        t.locate(8.0)
    with pytest.raises(ValueError):
        t.locate(-0.1)
    t.assign([2, 2], [0.0, 1.0])            # duplicate leaves: the later value stays
    assert t.node[t.width + 2] == 1.0 and t.total == 6.5


def test_importance_weights_known_answers():
    # replay.py:344-376: weights (u/p)^beta; normalised so that the largest is 1
    p = np.array([0.1, 0.2, 0.4])
    assert np.allclose(ro.is_weights(p, 0.2, 1.0, False), [2.0, 1.0, 0.5])
    assert np.allclose(ro.is_weights(p, 0.2, 0.5, True), np.sqrt([2.0, 1.0, 0.5]) / np.sqrt(2.0))
    assert np.allclose(ro.is_weights(p, 0.2, 0.0, False), 1.0)


@pytest.mark.skipif(not ref_shim.available(), reason="reference tree not on this machine")
@pytest.mark.parametrize("kind", ["mgsc_fifo", "mgsc_reservoir"])
def test_oracle_mgsc_matches_reference_fresh_seed(kind):
    from tests.golden.make_golden import reference_factory
    args = dict(seed=77, capacity=19, n_add=120, every=3, batch=4, meta=6)
    a = sc.pairs_mgsc(reference_factory(kind), **args)
    b = sc.pairs_mgsc(oracle_factory(kind), **args)
    for k in a:
        assert np.array_equal(a[k], b[k]), k


def test_reservoir_inclusion_statistics():
    # dqn_reservoir/run_reservoir_test.py:26-47 (fewer epochs, wider band)
    cap, n, epochs = 50, 200, 600
    hits = np.zeros(n)
    for epoch in range(epochs):
        rep = ro.ReservoirReplay(cap, sc.Pair(None, None), np.random.RandomState(epoch))
        for i in range(n):
            rep.add(sc.Pair(i, 0))
        for item in rep.items.values():
            hits[item.a] += 1
    freq = hits / epochs
    assert np.all(np.abs(freq - cap / n) < 0.09), (freq.min(), freq.max())


def test_fifo_ids_contiguous_window():
    # replay_test.py:129-147
    rep = ro.FifoReplay(10, sc.Pair(None, None), np.random.RandomState(0))
    for i in range(35):
        rep.add(sc.Pair(i, i))
        assert rep.size == min(i + 1, 10)
        assert list(rep.items.keys()) == list(range(max(0, i - 9), i + 1))


def test_uniform_ids_errors():
    # replay_test.py:51-54
    d = ro.UniformIds(np.random.RandomState(0))
    d.add([1, 2])
    with pytest.raises(IndexError, match="Cannot add ID"):
        d.add([2])
    with pytest.raises(IndexError, match="Cannot remove ID"):
        d.remove([5])


def test_stack_layout_trailing_zeros():
    # processors.py:497-504: A000, AB00, ABC0, ABCD, BCDE
    ts = synthetic.timestep_stream(np.random.default_rng(0), hw=(4, 4), short_every=0, min_len=6)
    obs = [next(ts).observation for _ in range(6)]
    assert not obs[0][..., 1:].any() and obs[0][..., 0].any()
    assert np.array_equal(obs[1][..., 0], obs[0][..., 0]) and not obs[1][..., 2:].any()
    assert np.array_equal(obs[4][..., :3], obs[3][..., 1:])
    assert np.array_equal(obs[5][..., :3], obs[4][..., 1:])
