"""The reference's own replay tests, run against the drop-in classes (host side, no GPU):
dqn_zoo/replay_test.py:39-175 (UniformDistributionTest, TransitionReplayTest) and the reservoir
inclusion statistics of dqn_zoo/dqn_reservoir/run_reservoir_test.py:26-47.  Items here are the
reference tests' `Pair(a, b)` of Python ints, which the product keeps in a host list exactly like
the reference does (frames go to HBM: tests/test_replay_gpu.py)."""
import collections
import copy

import numpy as np
import pytest

from dqn_mgsc_zoo_b200 import replay as replay_lib

Pair = collections.namedtuple('Pair', ['a', 'b'])


def assert_trees_equal(a, b):
    if isinstance(a, dict):
        assert set(a) == set(b)
        for k in a:
            assert_trees_equal(a[k], b[k])
    elif isinstance(a, (list, tuple)):
        assert len(a) == len(b)
        for x, y in zip(a, b):
            assert_trees_equal(x, y)
    elif isinstance(a, np.random.RandomState):
        sa, sb = a.get_state(), b.get_state()
        assert sa[0] == sb[0] and np.array_equal(sa[1], sb[1]) and sa[2:] == sb[2:]
    else:
        assert np.array_equal(a, b)


# ---- UniformDistributionTest (replay_test.py:39-100)
@pytest.fixture
def dist():
    return replay_lib.UniformDistribution(np.random.RandomState(seed=1))


def test_add_and_sample_one(dist):
    dist.add([2])
    assert dist.sample(1) == 2
    assert dist.check_valid()[0]


def test_adding_existing_id_raises_an_error(dist):
    dist.add([2, 5])
    with pytest.raises(IndexError, match='Cannot add ID'):
        dist.add([6, 5])


def test_remove(dist):
    dist.add([2, 5, 7])
    dist.remove([5])
    assert 5 not in dist.sample(100)
    assert dist.check_valid()[0]


def test_remove_final_id(dist):
    dist.add([2, 5, 7])
    dist.remove([7])
    assert dist.size == 2
    assert 7 not in dist.sample(100)
    assert dist.check_valid()[0]


def test_removing_nonexistent_id_raises_an_error(dist):
    dist.add([2, 5])
    with pytest.raises(IndexError, match='Cannot remove ID'):
        dist.remove([7])


def test_distribution_size(dist):
    dist.add([2, 5, 3])
    assert dist.size == 3
    dist.remove([5])
    assert dist.size == 2


def test_distribution_get_state_and_set_state(dist):
    dist.add([2, 5, 3, 8])
    dist.remove([5])
    assert dist.check_valid()[0]
    state = copy.deepcopy(dist.get_state())
    new_dist = replay_lib.UniformDistribution(np.random.RandomState(seed=1))
    new_dist.set_state(state)
    assert new_dist.check_valid()[0]
    assert new_dist.size == dist.size
    assert_trees_equal(state, new_dist.get_state())


# ---- TransitionReplayTest (replay_test.py:103-175)
CAPACITY = 10
ITEMS = [Pair(a=1, b=2), Pair(a=11, b=22), Pair(a=111, b=222), Pair(a=1111, b=2222)]


@pytest.fixture
def replay():
    r = replay_lib.TransitionReplay(capacity=CAPACITY, structure=Pair(a=None, b=None),
                                    random_state=np.random.RandomState(1))
    for item in ITEMS:
        r.add(item)
    return r


def test_size_and_capacity(replay):
    assert replay.size == len(ITEMS)
    assert replay.capacity == CAPACITY


def test_sample(replay):
    samples = replay.sample(2)
    assert samples.a.shape == (2,)
    assert np.array_equal(samples.b, 2 * samples.a)


def test_invariants():
    capacity, num_samples = 10, 3
    replay = replay_lib.TransitionReplay(capacity=capacity, structure=Pair(a=None, b=None),
                                         random_state=np.random.RandomState(1))
    for i in range(31):
        replay.add(Pair(a=i, b=2 * i))
        assert replay.size <= capacity
        if i > 2:
            assert replay.sample(num_samples).a.shape == (num_samples,)
        ids = list(replay.ids())
        assert len(ids) == min(i + 1, capacity)
        assert ids == sorted(ids)
        assert ids[0] + len(ids) - 1 == ids[-1]
        assert [p.a for p in replay.get(ids)] == ids          # FIFO: id i holds the i-th item


def test_get_state_and_set_state():
    make = lambda: replay_lib.TransitionReplay(capacity=CAPACITY, structure=Pair(a=None, b=None),
                                               random_state=np.random.RandomState(1))
    replay = make()
    for i in range(100):
        replay.add(Pair(a=i, b=i))
    replay.sample(10)
    assert replay.check_valid()[0]
    state = copy.deepcopy(replay.get_state())
    new_replay = make()
    new_replay.set_state(state)
    assert new_replay.check_valid()[0]
    assert new_replay.size == replay.size
    assert_trees_equal(state, new_replay.get_state())
    # both continue identically (same ids, same items) after the restore
    replay.add(Pair(a=100, b=100))
    new_replay.add(Pair(a=100, b=100))
    assert list(replay.ids()) == list(new_replay.ids())
    assert [p.a for p in replay.get(replay.ids())] == [p.a for p in new_replay.get(new_replay.ids())]


def test_empty_and_oversized_sample():
    replay = replay_lib.TransitionReplay(4, Pair(None, None), np.random.RandomState(0))
    with pytest.raises(ValueError):            # RandomState.randint(0): the reference raises the same
        replay.sample(1)
    replay.add(Pair(1, 2))
    assert replay.sample(5).a.tolist() == [1] * 5          # sampling is with replacement


# ---- reservoir inclusion statistics (dqn_reservoir/run_reservoir_test.py:26-47, fewer epochs)
def test_reservoir_keeps_every_item_with_probability_capacity_over_n():
    epochs, capacity, n, eps = 1500, 50, 200, 0.045
    structure = replay_lib.Transition(None, None, None, None, None)
    transitions = [replay_lib.Transition(i, i, i, i, i) for i in range(n)]
    counts = collections.Counter()
    for epoch in range(epochs):
        replay = replay_lib.ReservoirTransitionReplay(capacity, structure, np.random.RandomState(epoch), None, None)
        for tr in transitions:
            replay.add(tr)
        assert replay.size == capacity
        for tr in replay.get(replay.ids()):
            counts[tr.a_tm1] += 1
    expected = capacity / n
    freq = np.array([counts[i] / epochs for i in range(n)])
    assert np.all(np.abs(freq - expected) <= eps), (freq.min(), freq.max())


def test_transition_accumulator_pairs_consecutive_timesteps():
    # replay.py:1163-1197 (the agents' accumulator; the n-step one is out of scope)
    from dqn_mgsc_zoo_b200 import synthetic
    acc = replay_lib.TransitionAccumulator()
    obs = [np.full((2, 2), i, np.uint8) for i in range(4)]
    first = synthetic.TimeStep(0, None, None, obs[0])
    assert list(acc.step(first, 3)) == []
    mid = synthetic.TimeStep(1, 0.5, 0.9, obs[1])
    (t,) = list(acc.step(mid, 4))
    assert t.s_tm1 is obs[0] and t.s_t is obs[1] and t.a_tm1 == 3 and t.r_t == 0.5 and t.discount_t == 0.9
    last = synthetic.TimeStep(2, 1.0, 0.0, obs[2])
    (t,) = list(acc.step(last, 5))
    assert t.s_tm1 is obs[1] and t.a_tm1 == 4 and t.discount_t == 0.0
    acc.reset()
    with pytest.raises(ValueError, match='Expected FIRST timestep'):
        list(acc.step(mid, 0))


def test_uniform_distribution_id_mirror_tracks_the_list():
    """sample() takes from an int64 mirror of `_ids`; adds, swap-removes and set_state must keep the
    mirror equal to the list, and the draw must be the reference's (positions -> ids)."""
    import numpy as np
    from dqn_mgsc_zoo_b200 import replay as rp
    d = rp.UniformDistribution(np.random.RandomState(4))
    ref_rng = np.random.RandomState(4)
    rng = np.random.default_rng(0)
    next_id = 0
    for step in range(5000):
        if d.size < 3 or rng.random() < 0.6:
            d.add([next_id]); next_id += 1
        else:
            d.remove([int(d._ids[int(rng.integers(d.size))])])
        if step % 97 == 0:
            assert np.array_equal(d._ids_np[:d.size], np.asarray(d._ids, dtype=np.int64))
            got = d.sample(16)
            want = np.asarray([d._ids[p] for p in ref_rng.randint(d.size, size=16)], dtype=np.int64)
            assert got.dtype == np.int64 and np.array_equal(got, want)
    e = rp.UniformDistribution(np.random.RandomState(4))
    e.set_state(d.get_state())
    assert np.array_equal(e._ids_np[:e.size], np.asarray(d._ids, dtype=np.int64))
    assert d.check_valid()[0]


# ---- golden fixtures made by the unmodified reference, product classes on generic (host) items
@pytest.mark.parametrize("name", ["fifo_pairs", "fifo_pairs_tiny", "reservoir_pairs", "gen_fifo_pairs",
                                  "gen_reservoir_pairs"])
def test_product_pairs_match_reference_golden(name):
    from dqn_mgsc_zoo_b200 import replay_circular as rc
    from tests import replay_scenarios as sc
    from tests.test_oracle_replay import check_against_golden
    spec = sc.SCENARIOS[name]
    cls = {"fifo": replay_lib.TransitionReplay, "reservoir": replay_lib.ReservoirTransitionReplay,
           "gen_fifo": rc.TransitionReplay, "gen_reservoir": rc.ReservoirTransitionReplay}[spec["kind"]]
    got = sc.pairs_uniform(lambda capacity, structure, rng: cls(capacity, structure, rng), **spec["args"])
    check_against_golden(name, got)


def test_replay_circular_module_surface_of_the_mgsc_runner():
    """dqn_mgsc_batched/run_atari.py:185-232 imports replay_circular as replay_lib and touches exactly these."""
    from dqn_mgsc_zoo_b200 import replay_circular as rc
    for name in ("Transition", "TransitionAccumulator", "NStepTransitionAccumulator", "compress_array",
                 "uncompress_array", "CircularBuffer", "CircularLogitBuffer", "UniformDistribution",
                 "TransitionReplay", "ReservoirTransitionReplay", "MGSCReservoirDistribution",
                 "MGSCFiFoTransitionReplay", "MGSCReservoirTransitionReplay"):
        assert hasattr(rc, name), name
    a = np.arange(84 * 84 * 4, dtype=np.uint8).reshape(84, 84, 4)
    t = rc.Transition(s_tm1=a, a_tm1=1, r_t=0.0, discount_t=0.99, s_t=a + 1)
    enc = t._replace(s_tm1=rc.compress_array(t.s_tm1), s_t=rc.compress_array(t.s_t))       # the runner's encoder
    dec = enc._replace(s_tm1=rc.uncompress_array(enc.s_tm1), s_t=rc.uncompress_array(enc.s_t))
    assert np.array_equal(dec.s_tm1, t.s_tm1) and np.array_equal(dec.s_t, t.s_t)
    buf = rc.CircularBuffer(3)
    with pytest.raises(BufferError):
        buf.popleft()
    for i in range(3):
        buf.add(i)
    with pytest.raises(BufferError):
        buf.add(9)
    assert buf.popleft() == 0 and buf[0] == 1 and len(buf) == 2
    buf.add(7)
    assert [buf[k] for k in range(3)] == [1, 2, 7]
    st = buf.get_state()
    other = rc.CircularBuffer(3)
    other.set_state(st)
    assert [other[k] for k in range(3)] == [1, 2, 7]


def test_n_step_accumulator_matches_reference_semantics():
    """replay_test.py:177-360 (NStepTransitionAccumulatorTest): n-step returns and episode-end flush."""
    import collections
    TS = collections.namedtuple("TS", "step_type reward discount observation")
    TS.first = lambda self: self.step_type == 0
    TS.last = lambda self: self.step_type == 2
    acc = replay_lib.NStepTransitionAccumulator(3)
    steps = [TS(0, None, None, 0)] + [TS(1, float(k), 0.5, k) for k in range(1, 5)] + [TS(2, 5.0, 0.0, 5)]
    out = []
    for k, ts in enumerate(steps):
        out.append(list(acc.step(ts, a_t=k)))
    assert [len(o) for o in out] == [0, 0, 0, 1, 1, 3]
    first = out[3][0]
    assert (first.s_tm1, first.a_tm1, first.s_t) == (0, 0, 3)
    assert first.r_t == 1.0 + 0.5 * 2.0 + 0.25 * 3.0 and first.discount_t == 0.125
    tail = out[5]
    assert [(t.s_tm1, t.s_t) for t in tail] == [(2, 5), (3, 5), (4, 5)]
    assert tail[-1].r_t == 5.0 and tail[-1].discount_t == 0.0
    with pytest.raises(ValueError, match="Expected FIRST"):
        replay_lib.NStepTransitionAccumulator(2).step(TS(1, 0.0, 1.0, 0), 0).__next__()


def test_inlined_single_id_bookkeeping_equals_the_distribution_methods():
    """TransitionReplay.add writes remove([oldest]) + add([new]) out for the stock UniformDistribution; any other
    distribution type (here: a trivial subclass) goes through the methods.  Both must leave identical state and
    draw identical samples, through the fill, the first wrap and many wraps."""
    class Routed(replay_lib.UniformDistribution):
        pass

    class RoutedReplay(replay_lib.TransitionReplay):
        _DISTRIBUTION = Routed

    capacity = 37
    fast = replay_lib.TransitionReplay(capacity, Pair(a=None, b=None), np.random.RandomState(11))
    slow = RoutedReplay(capacity, Pair(a=None, b=None), np.random.RandomState(11))
    assert type(fast._distribution) is replay_lib.UniformDistribution and type(slow._distribution) is Routed
    for i in range(5 * capacity + 3):
        fast.add(Pair(a=i, b=2 * i))
        slow.add(Pair(a=i, b=2 * i))
        if i % 7 == 0:
            df, ds = fast._distribution, slow._distribution
            assert df._ids == ds._ids and df._id_to_index == ds._id_to_index
            assert np.array_equal(df._ids_np[:df.size], ds._ids_np[:ds.size])
            assert list(fast.ids()) == list(slow.ids()) and fast._t == slow._t
            if i > 3:
                assert np.array_equal(fast.sample_slots(5), slow.sample_slots(5))
    assert fast.check_valid()[0] and slow.check_valid()[0]
