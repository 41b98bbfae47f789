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
