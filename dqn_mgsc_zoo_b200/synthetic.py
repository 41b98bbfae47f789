"""Synthetic Atari-shaped workload: processed timesteps with the
reference's 4-frame stack layout, for tests, smoke() and bench.py.

The ALE emulator and `processors.atari` stay on the host and are outside the
optimisation target; what matters to the replay store is the *layout* of the
stacks the processor emits (processors.py:489-505): frames oldest-first along
the last axis, zero planes TRAILING at episode start:
    A000, AB00, ABC0, ABCD, BCDE, ...
"""
import collections

import numpy as np

FRAME_HW = (84, 84)
STACK = 4


class TimeStep(collections.namedtuple("TimeStep", "step_type reward discount observation")):
    """Minimal stand-in for dm_env.TimeStep (FIRST=0, MID=1, LAST=2)."""
    __slots__ = ()

    def first(self):
        return self.step_type == 0

    def mid(self):
        return self.step_type == 1

    def last(self):
        return self.step_type == 2


def stack_frames(frames):
    """Deque(4) -> list -> trailing_zero_pad(4) -> np.stack(axis=-1)."""
    frames = list(frames)[-STACK:]
    pad = [np.zeros_like(frames[0])] * (STACK - len(frames))
    return np.stack(frames + pad, axis=-1)


def episode_lengths(rng, short_every=97):
    """Episode lengths in agent steps; every `short_every`-th episode is 1..4 long."""
    n = 0
    while True:
        n += 1
        if short_every and n % short_every == 0:
            yield int(1 + (n // short_every) % 4)
        else:
            yield int(rng.integers(50, 2000))


def timestep_stream(rng, hw=FRAME_HW, short_every=97, max_len=None, min_len=None):
    """Yields processed TimeSteps forever: FIRST (reward/discount None) then MID.. LAST."""
    for length in episode_lengths(rng, short_every):
        if max_len is not None:
            length = min(length, max_len)
        if min_len is not None:
            length = max(length, min_len)
        window = collections.deque(maxlen=STACK)
        for k in range(length + 1):
            window.append(rng.integers(0, 256, size=hw, dtype=np.uint8))
            obs = stack_frames(window)
            if k == 0:
                yield TimeStep(0, None, None, obs)
                continue
            reward = float(rng.choice([-1.0, 0.0, 1.0], p=[0.05, 0.90, 0.05]))
            last = k == length
            discount = 0.0 if (last or rng.random() < 0.01) else 0.99
            yield TimeStep(2 if last else 1, reward, discount, obs)


def transition_stream(rng, num_actions, accumulator, **kw):
    """Feeds the timestep stream through a TransitionAccumulator; yields transitions."""
    for ts in timestep_stream(rng, **kw):
        action = int(rng.integers(0, num_actions))
        for tr in accumulator.step(ts, action):
            yield tr


def random_batch(rng, batch, num_actions, hw=FRAME_HW):
    """Independent random (s_tm1, a, r, discount, s_t) arrays for learner tests."""
    s_tm1 = rng.integers(0, 256, size=(batch,) + tuple(hw) + (STACK,), dtype=np.uint8)
    s_t = rng.integers(0, 256, size=(batch,) + tuple(hw) + (STACK,), dtype=np.uint8)
    a = rng.integers(0, num_actions, size=batch).astype(np.int32)
    r = rng.choice([-1.0, 0.0, 1.0], size=batch, p=[0.05, 0.90, 0.05]).astype(np.float32)
    d = np.where(rng.random(batch) < 0.02, 0.0, 0.99).astype(np.float32)
    return s_tm1, a, r, d, s_t
