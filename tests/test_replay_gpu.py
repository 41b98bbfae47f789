"""GPU parity of the HBM replay store, through the C ABI, bit-exact against (1) the golden
fixtures produced by the unmodified reference and (2) the oracle on larger seeded streams."""
import numpy as np
import pytest

from dqn_mgsc_zoo_b200 import replay as rp
from dqn_mgsc_zoo_b200 import synthetic
from oracle import replay_oracle as ro
from tests import replay_scenarios as sc
from tests.test_oracle_replay import check_against_golden

pytestmark = pytest.mark.gpu

UNIFORM = [n for n, s in sc.SCENARIOS.items() if s["kind"] in ("fifo", "reservoir", "gen_fifo", "gen_reservoir")]


def product_factory(kind):
    from dqn_mgsc_zoo_b200 import replay_circular as rc
    cls = {"fifo": rp.TransitionReplay, "reservoir": rp.ReservoirTransitionReplay,
           "gen_fifo": rc.TransitionReplay, "gen_reservoir": rc.ReservoirTransitionReplay}[kind]
    return lambda capacity, structure, rng: cls(capacity, structure, rng)


@pytest.mark.parametrize("name", UNIFORM)
def test_product_matches_reference_golden(name):
    spec = sc.SCENARIOS[name]
    fn = getattr(sc, spec["fn"])
    if spec["fn"] == "frames_uniform":
        got = fn(product_factory(spec["kind"]), rp.TransitionAccumulator(),
                 rp.Transition(None, None, None, None, None), **spec["args"])
    else:
        got = fn(product_factory(spec["kind"]), **spec["args"])
    check_against_golden(name, got)


@pytest.mark.parametrize("kind,hw", [("fifo", (84, 84)), ("reservoir", (84, 84)), ("fifo", (10, 6))])
def test_product_matches_oracle_full_arrays(kind, hw):
    cap, n_add, batch = 300, 1500, 32
    o_cls = {"fifo": ro.FifoReplay, "reservoir": ro.ReservoirReplay}[kind]
    o = o_cls(cap, ro.Transition(None, None, None, None, None), np.random.RandomState(5))
    p = product_factory(kind)(cap, rp.Transition(None, None, None, None, None), np.random.RandomState(5))
    so = synthetic.transition_stream(np.random.default_rng(9), 6, ro.PairAccumulator(), hw=hw,
                                     short_every=7, max_len=40)
    sp = synthetic.transition_stream(np.random.default_rng(9), 6, rp.TransitionAccumulator(), hw=hw,
                                     short_every=7, max_len=40)
    for i in range(n_add):
        o.add(next(so))
        p.add(next(sp))
        if i % 37 == 36:
            a, b = o.sample(batch), p.sample(batch)
            for x, y, name in zip(a, b, a._fields):
                assert x.dtype == y.dtype and x.shape == y.shape, name
                assert np.array_equal(x, y), (i, name)
    st = p.frame_store.stats()
    assert st["puts"] >= cap
    if hw == (84, 84) and kind == "fifo":
        # chained transitions upload ~1 plane each instead of 8
        assert st["planes_uploaded"] < 2.0 * st["puts"], st


def test_get_and_state_roundtrip():
    p = rp.TransitionReplay(20, rp.Transition(None, None, None, None, None), np.random.RandomState(1))
    stream = synthetic.transition_stream(np.random.default_rng(2), 4, rp.TransitionAccumulator(),
                                         hw=(12, 8), max_len=9)
    items = [next(stream) for _ in range(33)]
    for it in items:
        p.add(it)
    assert list(p.ids()) == list(range(13, 33))
    for got, want in zip(p.get([13, 20, 32]), (items[13], items[20], items[32])):
        for x, y in zip(got, want):
            assert np.array_equal(x, y)
    state = p.get_state()
    q = rp.TransitionReplay(20, rp.Transition(None, None, None, None, None), np.random.RandomState(1))
    q.set_state(state)
    p._random_state.seed(3)
    q._random_state.seed(3)
    a, b = p.sample(16), q.sample(16)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)
    assert q.check_valid() == (True, '')


def test_gather_split_settings_agree():
    p = rp.TransitionReplay(64, rp.Transition(None, None, None, None, None), np.random.RandomState(1))
    stream = synthetic.transition_stream(np.random.default_rng(4), 4, rp.TransitionAccumulator(), max_len=11)
    for _ in range(64):
        p.add(next(stream))
    slots = np.arange(64)
    base = p.frame_store.gather_host(slots)
    for splits in (1, 2, 3, 7, 16):
        p.frame_store.set_gather_splits(splits)
        for x, y in zip(base, p.frame_store.gather_host(slots)):
            assert np.array_equal(x, y), splits


def test_commit_order_large_then_small_batches_same_slots():
    """A >128 KB staged batch is committed on the caller's stream, a small one on the store's side
    stream; with capacity 4 both write the same slots, so the commits must land in issue order
    (the older SlotRecord could land last).  Non-chained puts: 8 new planes each."""
    cap = 4
    rng = np.random.default_rng(21)
    p = rp.TransitionReplay(cap, rp.Transition(None, None, None, None, None), np.random.RandomState(0))
    o = ro.FifoReplay(cap, ro.Transition(None, None, None, None, None), np.random.RandomState(0))

    def item():
        s0 = rng.integers(1, 256, size=(84, 84, 4), dtype=np.uint8)
        s1 = rng.integers(1, 256, size=(84, 84, 4), dtype=np.uint8)
        return (s0, int(rng.integers(6)), float(rng.integers(-1, 2)), 0.99, s1)

    for rounds in range(6):
        for n in (70, 3, 2, 33, 1):      # big staged batches (copied) followed by small (zero-copy) ones
            for _ in range(n):
                it = item()
                p.add(rp.Transition(*it))
                o.add(ro.Transition(*it))
            a, b = o.sample(8), p.sample(8)
            for x, y, name in zip(a, b, a._fields):
                assert np.array_equal(x, y), (rounds, n, name)


# ---------------------------------------------------------------- prioritized replay (N3)
PER = [n for n, s in sc.SCENARIOS.items() if s["kind"] in ("per", "gen_per")]


def per_factory(kind):
    from dqn_mgsc_zoo_b200 import replay_circular as rc
    cls = (rp if kind == "per" else rc).PrioritizedTransitionReplay
    return lambda capacity, structure, alpha, beta, usp, normalize, rng: cls(
        capacity, structure, alpha, beta, usp, normalize, rng)


@pytest.mark.parametrize("name", PER)
def test_prioritized_replay_matches_reference_golden(name):
    """Device sum tree: sampled ids bit-exact, importance weights bit-exact (float64 on the host from bit-exact
    probabilities), the caller's next random draw identical - against fixtures made by the unmodified reference
    PrioritizedTransitionReplay (replay.py:1046-1160 / replay_circular.py:1315-1430)."""
    spec = sc.SCENARIOS[name]
    got = sc.pairs_prioritized(per_factory(spec["kind"]), **spec["args"])
    check_against_golden(name, got)


def test_sum_tree_matches_oracle_tree_and_reference_api():
    """SumTreeTest (replay_test.py:820-1045): set / get / query / root / resize / set_all / state, node for node
    against the oracle tree (same float64 additions in tree order => identical bits)."""
    rng = np.random.default_rng(3)
    tree = rp.SumTree()
    assert tree.size == 0 and np.isnan(tree.root())
    tree.resize(37)
    assert tree.size == 37 and tree.capacity == 64 and tree.root() == 0.0
    o = ro.PrefixTree(37)
    for _ in range(40):
        idx = rng.integers(0, 37, size=int(rng.integers(1, 9)))
        val = np.where(rng.random(idx.size) < 0.2, 0.0, rng.random(idx.size) * 3.0)
        tree.set(idx, val)
        o.assign(idx, val)
        assert tree.root() == o.total
    st = tree.get_state()
    assert st['first_leaf'] == 64 and np.array_equal(st['storage'], o.node)
    targets = rng.random(200) * o.total
    assert tree.query(targets) == [o.locate(t) for t in targets]
    assert np.array_equal(tree.get(np.arange(37)), o.node[64:64 + 37])
    assert tree.check_valid() == (True, '')
    with pytest.raises(ValueError, match='Require 0 <= target'):
        tree.query([o.total])
    with pytest.raises(ValueError, match='finite and positive'):
        tree.set([0], [-1.0])
    with pytest.raises(IndexError):
        tree.get([37])
    # resize: grow within capacity, grow beyond it (values kept), shrink (tail zeroed)
    tree.resize(50)
    assert tree.size == 50 and tree.capacity == 64 and tree.root() == o.total
    tree.resize(100)
    assert tree.capacity == 128 and np.array_equal(tree.values[:37], o.node[64:64 + 37]) and tree.check_valid()[0]
    tree.resize(10)
    assert tree.size == 10 and tree.check_valid() == (True, '') and not tree.get_state()['storage'][tree.capacity + 10:].any()
    tree.set_all(np.arange(1, 6, dtype=np.float64))
    assert tree.size == 5 and tree.root() == 15.0 and tree.query([0.5, 1.0, 14.9]) == [0, 1, 4]
    other = rp.SumTree()
    other.set_state(tree.get_state())
    assert other.root() == 15.0 and other.check_valid() == (True, '')


def test_prioritized_replay_frames_match_oracle():
    """Atari-shaped items at capacity 300 (FIFO wrap, non power-of-two), 84x84: batches bit-exact, ids exact,
    weights exact, same RandomState consumption, state round trip."""
    cap, n_add, batch = 300, 1200, 32
    beta = lambda t: min(1.0, 0.4 + t * 5e-4)
    o = ro.PrioritizedReplay(cap, ro.Transition(None, None, None, None, None), 0.6, beta, 1e-3, True,
                             np.random.RandomState(8))
    p = rp.PrioritizedTransitionReplay(cap, rp.Transition(None, None, None, None, None), 0.6, beta, 1e-3, True,
                                       np.random.RandomState(8))
    so = synthetic.transition_stream(np.random.default_rng(9), 6, ro.PairAccumulator(), short_every=7, max_len=40)
    sp = synthetic.transition_stream(np.random.default_rng(9), 6, rp.TransitionAccumulator(), short_every=7, max_len=40)
    rng = np.random.default_rng(1)
    max_pri, checks = 1.0, 0
    for i in range(n_add):
        o.add(next(so), max_pri)
        p.add(next(sp), max_pri)
        if i % 23 == 22 and o.size >= batch:
            (a, ia, wa), (b, ib, wb) = o.sample(batch), p.sample(batch)
            assert np.array_equal(ia, ib) and np.array_equal(wa, wb), i
            for x, y, f in zip(a, b, a._fields):
                assert np.array_equal(x, y), (i, f)
            pri = np.abs(rng.normal(size=batch)).astype(np.float32)
            max_pri = float(max(max_pri, pri.max()))
            o.update_priorities(ia, pri)
            p.update_priorities(ib, pri)
            checks += 1
    assert checks >= 40 and p.size == cap and p.check_valid() == (True, '')
    assert p._random_state.randint(1 << 30) == o.rs.randint(1 << 30)
    q = rp.PrioritizedTransitionReplay(cap, rp.Transition(None, None, None, None, None), 0.6, beta, 1e-3, True,
                                       np.random.RandomState(8))
    q.set_state(p.get_state())
    p._random_state.seed(4)
    q._random_state.seed(4)
    (a, ia, wa), (b, ib, wb) = p.sample(16), q.sample(16)
    assert np.array_equal(ia, ib) and np.array_equal(wa, wb)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)


# ---------------------------------------------------------------- checkpoint (N2)
def _same_samples(p, q, seed, n=24, mgsc=False):
    for rep in (p, q):
        rs = rep._distribution._rng_state if mgsc else rep._random_state
        if mgsc:
            rs.bit_generator.state = np.random.default_rng(seed).bit_generator.state
        else:
            rs.seed(seed)
    a, b = p.sample(n), q.sample(n)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)


@pytest.mark.parametrize("kind", ["fifo", "reservoir", "mgsc_fifo", "mgsc_reservoir"])
def test_snapshot_state_round_trip_resumes_identically(kind):
    """get_state() exports the HBM store as one blob behind a lazy reference-shaped view; after a pickle round
    trip set_state() imports it in one pass.  The restored replay samples identically, keeps accepting adds with
    the same dedup/chaining state (ref-counts, free list, last stack) and still matches the live one afterwards."""
    import pickle
    from dqn_mgsc_zoo_b200 import replay_circular as rc
    cap = 200
    structure = rp.Transition(None, None, None, None, None)
    mgsc = kind.startswith("mgsc")
    cls = {"fifo": rp.TransitionReplay, "reservoir": rp.ReservoirTransitionReplay,
           "mgsc_fifo": rc.MGSCFiFoTransitionReplay, "mgsc_reservoir": rc.MGSCReservoirTransitionReplay}[kind]
    mk = lambda seed: cls(cap, structure, np.random.default_rng(seed) if mgsc else np.random.RandomState(seed))
    p = mk(1)
    stream = synthetic.transition_stream(np.random.default_rng(5), 6, rp.TransitionAccumulator(), short_every=5, max_len=30)
    items = [next(stream) for _ in range(700)]
    for it in items[:450]:
        p.add(it)
    state = pickle.loads(pickle.dumps(p.get_state(), protocol=4))
    q = mk(2)
    q.set_state(state)
    assert q.size == p.size
    if kind in ("fifo", "reservoir"):
        # the lazy view is reference-shaped: (id, item) pairs that materialise to the stored bytes
        i0, item0 = state['storage'][3]
        want = next(p.get([i0]))
        assert all(np.array_equal(x, y) for x, y in zip(item0, want))
    if not mgsc:
        q._random_state.set_state(p._random_state.get_state())     # reservoir decisions must continue identically
    else:
        q_rng = q._distribution._rng_state if kind == "mgsc_fifo" else q._random_state
        p_rng = p._distribution._rng_state if kind == "mgsc_fifo" else p._random_state
        q_rng.bit_generator.state = p_rng.bit_generator.state
    _same_samples(p, q, 11, mgsc=mgsc)
    for it in items[450:]:
        p.add(it)
        q.add(it)
    _same_samples(p, q, 12, mgsc=mgsc)
    assert p.frame_store.stats()["planes_in_use"] == q.frame_store.stats()["planes_in_use"]


def test_pipelined_gather_equals_per_transition_gather():
    """Large launches give every gather CTA 5 plane buffers (6 CTAs per SM) and run rows with more unique planes in
    two phases; small ones keep 8 buffers.  Same bytes either way and for any CTA split, including early-episode
    stacks (zero planes), reservoir slots (8 unique planes), repeated slots and odd row counts.  (DQNB_GATHER_PIPE_MIN
    selects the persistent pipelined variant, measured slower and off by default.)"""
    p = rp.ReservoirTransitionReplay(400, rp.Transition(None, None, None, None, None), np.random.RandomState(1))
    stream = synthetic.transition_stream(np.random.default_rng(4), 4, rp.TransitionAccumulator(), short_every=3, max_len=9)
    for _ in range(900):                     # reservoir replacements: slots with no temporal neighbours (8 unique planes)
        p.add(next(stream))
    store = p.frame_store
    small = store.gather_host(np.arange(200))          # n < 256: 8 plane buffers per CTA, single phase
    big = store.gather_host(np.concatenate([np.arange(200), np.arange(200)]))   # n >= 256: 5 buffers, two-phase rows
    for x, y in zip(small, big):
        assert np.array_equal(x, y[:200]) and np.array_equal(x, y[200:])
    for n in (1024, 1500, 4099):
        slots = np.random.default_rng(n).integers(0, 400, size=n)
        store.set_gather_splits(0)
        fast = store.gather_host(slots)
        store.set_gather_splits(2)
        base = store.gather_host(slots)
        for x, y in zip(base, fast):
            assert np.array_equal(x, y), n
    store.set_gather_splits(0)
    o = ro.ReservoirReplay(400, ro.Transition(None, None, None, None, None), np.random.RandomState(1))
    so = synthetic.transition_stream(np.random.default_rng(4), 4, ro.PairAccumulator(), short_every=3, max_len=9)
    for _ in range(900):
        o.add(next(so))
    ids = np.random.default_rng(7).integers(0, 400, size=2048)
    want = o.gather(ids)
    got = store.gather_host(ids)
    assert np.array_equal(got[0], want.s_tm1) and np.array_equal(got[4], want.s_t)
    assert np.array_equal(got[1], np.asarray(want.a_tm1, np.int32))
