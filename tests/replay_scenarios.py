"""Deterministic replay call sequences, driven through the PUBLIC replay API only
(add / sample / size / batch_of_ids_transitions_and_logits / update_priorities),
so the same scenario runs against the unmodified reference classes (to make the
golden fixtures), the oracle restatement and the CUDA-backed product classes."""
import collections
import zlib

import numpy as np

from dqn_mgsc_zoo_b200 import synthetic

Pair = collections.namedtuple("Pair", "a b")


def _crc(arr):
    return zlib.crc32(np.ascontiguousarray(arr).tobytes())


def pairs_uniform(make, seed, capacity, n_add, sample_every, batch, generator=False):
    """FIFO / reservoir with a RandomState (replay.py) or a Generator (the replay_circular.py twins):
    record every sampled `a` column."""
    rs = np.random.default_rng(seed) if generator else np.random.RandomState(seed)
    rep = make(capacity, Pair(None, None), rs)
    drawn, sizes = [], []
    for i in range(n_add):
        rep.add(Pair(i, 2 * i + 1))
        if i % sample_every == sample_every - 1:
            s = rep.sample(batch)
            assert np.array_equal(s.b, 2 * s.a + 1)
            drawn.append(np.asarray(s.a, dtype=np.int64))
            sizes.append(rep.size)
    # the caller shares the RandomState with the replay: its next draw is part of the contract
    return {"drawn": np.stack(drawn), "sizes": np.array(sizes),
            "next_draw": rs.integers(1 << 30) if generator else rs.randint(1 << 30)}


def pairs_mgsc(make, seed, capacity, n_add, every, batch, meta):
    """MGSC replays with a Generator: weighted samples, meta batches, priority writes."""
    rng = np.random.default_rng(seed)
    rep = make(capacity, Pair(None, None), rng)
    drawn, meta_idx, meta_logits = [], [], []
    for i in range(n_add):
        rep.add(Pair(i, 2 * i + 1))
        if i % every == every - 1 and rep.size >= max(batch, meta):
            idx, tr, lg = rep.batch_of_ids_transitions_and_logits(meta)
            meta_idx.append(np.asarray(idx, dtype=np.int64))
            meta_logits.append(np.asarray(lg, dtype=np.float32))
            new = np.asarray(lg, dtype=np.float32) + np.float32(0.25) * np.cos(np.asarray(tr.a)).astype(np.float32)
            rep.update_priorities(idx, new)
            drawn.append(np.asarray(rep.sample(batch).a, dtype=np.int64))
    return {"drawn": np.stack(drawn), "meta_idx": np.stack(meta_idx),
            "meta_logits": np.stack(meta_logits), "next_draw": rng.integers(1 << 30)}


def pairs_prioritized(make, seed, capacity, n_add, every, batch, alpha=0.6, usp=0.1, normalize=True,
                      generator=False, zero_until=0):
    """Prioritized replay: adds with priorities (the first `zero_until` all zero: the uniform fall-back branch of
    sample, which consumes one uniform draw less), samples, priority updates from a td-like signal."""
    rs = np.random.default_rng(seed) if generator else np.random.RandomState(seed)
    beta = lambda t: min(1.0, 0.4 + 0.6 * t / float(n_add))
    rep = make(capacity, Pair(None, None), alpha, beta, usp, normalize, rs)
    ids_log, w_log, a_log = [], [], []
    max_pri = 1.0
    for i in range(n_add):
        rep.add(Pair(i, 2 * i + 1), 0.0 if i < zero_until else max_pri)
        if i % every == every - 1 and rep.size >= batch:
            tr, ids, w = rep.sample(batch)
            assert np.array_equal(tr.b, 2 * tr.a + 1)
            ids_log.append(np.asarray(ids, dtype=np.int64))
            w_log.append(np.asarray(w, dtype=np.float64))
            a_log.append(np.asarray(tr.a, dtype=np.int64))
            pri = np.abs(np.sin(np.asarray(tr.a) * 0.37 + i).astype(np.float32)) * np.float32(2.0)   # |td| as float32
            if i % (3 * every) == every - 1:
                pri[0] = 0.0                       # an exact zero priority: 0 ** alpha stays 0
            max_pri = float(max(max_pri, pri.max()))
            rep.update_priorities(ids, pri)
    return {"ids": np.stack(ids_log), "weights": np.stack(w_log), "a": np.stack(a_log),
            "next_draw": rs.integers(1 << 30) if generator else rs.randint(1 << 30)}


def frames_uniform(make, accumulator, structure, seed, capacity, n_add, sample_every, batch, num_actions=6,
                   hw=(84, 84), short_every=5, max_len=23):
    """Atari-shaped transitions (short episodes => many early-episode stacks)."""
    rs = np.random.RandomState(seed)
    rep = make(capacity, structure, rs)
    stream = synthetic.transition_stream(np.random.default_rng(seed + 1), num_actions, accumulator,
                                         hw=hw, short_every=short_every, max_len=max_len)
    crcs, acts, rews, discs = [], [], [], []
    for i in range(n_add):
        rep.add(next(stream))
        if i % sample_every == sample_every - 1:
            s = rep.sample(batch)
            crcs.append([_crc(s.s_tm1), _crc(s.s_t)])
            acts.append(np.asarray(s.a_tm1, dtype=np.int64))
            rews.append(np.asarray(s.r_t, dtype=np.float64))
            discs.append(np.asarray(s.discount_t, dtype=np.float64))
    return {"crc": np.array(crcs, dtype=np.int64), "a": np.stack(acts), "r": np.stack(rews),
            "d": np.stack(discs), "shape": np.array(s.s_tm1.shape), "dtype": str(s.s_tm1.dtype)}


SCENARIOS = {
    "fifo_pairs": dict(kind="fifo", fn="pairs_uniform", args=dict(seed=7, capacity=37, n_add=180, sample_every=3, batch=8)),
    "fifo_pairs_tiny": dict(kind="fifo", fn="pairs_uniform", args=dict(seed=1, capacity=4, n_add=40, sample_every=1, batch=5)),
    "reservoir_pairs": dict(kind="reservoir", fn="pairs_uniform", args=dict(seed=11, capacity=50, n_add=600, sample_every=4, batch=8)),
    "gen_fifo_pairs": dict(kind="gen_fifo", fn="pairs_uniform", args=dict(seed=19, capacity=23, n_add=150, sample_every=3, batch=8, generator=True)),
    "gen_reservoir_pairs": dict(kind="gen_reservoir", fn="pairs_uniform", args=dict(seed=23, capacity=40, n_add=500, sample_every=4, batch=8, generator=True)),
    "per_pairs": dict(kind="per", fn="pairs_prioritized", args=dict(seed=29, capacity=37, n_add=260, every=3, batch=8)),
    "per_pairs_zero_start": dict(kind="per", fn="pairs_prioritized", args=dict(seed=31, capacity=16, n_add=90, every=2, batch=6, usp=0.25, normalize=False, zero_until=30)),
    "gen_per_pairs": dict(kind="gen_per", fn="pairs_prioritized", args=dict(seed=37, capacity=50, n_add=300, every=4, batch=8, generator=True)),
    "mgsc_fifo_pairs": dict(kind="mgsc_fifo", fn="pairs_mgsc", args=dict(seed=3, capacity=24, n_add=90, every=2, batch=6, meta=5)),
    "mgsc_reservoir_pairs": dict(kind="mgsc_reservoir", fn="pairs_mgsc", args=dict(seed=5, capacity=24, n_add=140, every=2, batch=6, meta=5)),
    "fifo_frames": dict(kind="fifo", fn="frames_uniform", args=dict(seed=13, capacity=64, n_add=260, sample_every=10, batch=16)),
    "reservoir_frames": dict(kind="reservoir", fn="frames_uniform", args=dict(seed=17, capacity=48, n_add=300, sample_every=10, batch=16)),
}
