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


class SyntheticEnvironment:
    """dm_env-style environment emitting random 84x84 uint8 frames (stands in for
    gym_atari.GymAtari, which stays on the host and outside the path): `reset()` -> FIRST
    timestep, `step(action)` -> MID ... LAST after `episode_length(rng)` steps."""

    def __init__(self, rng, num_actions=6, hw=FRAME_HW, min_len=5, max_len=60):
        self._rng, self._num_actions, self._hw = rng, num_actions, hw
        self._min_len, self._max_len = min_len, max_len
        self._left = 0

    def _frame(self):
        return self._rng.integers(0, 256, size=self._hw, dtype=np.uint8)

    def reset(self):
        self._left = int(self._rng.integers(self._min_len, self._max_len + 1))
        return TimeStep(0, None, None, self._frame())

    def step(self, action):
        if not 0 <= int(action) < self._num_actions:
            raise ValueError("action %r out of range" % (action,))
        self._left -= 1
        last = self._left <= 0
        reward = float(self._rng.choice([-1.0, 0.0, 1.0], p=[0.05, 0.90, 0.05]))
        return TimeStep(2 if last else 1, reward, 0.0 if last else 1.0, self._frame())


class StackingPreprocessor:
    """Minimal stand-in for `processors.atari` (processors.py:421-508) on frames that are
    already 84x84 grayscale: action repeat (returns None on skipped frames, summing rewards),
    4-frame stack with TRAILING zero padding, discount x0.99.  Has `reset()` like the
    reference processor so `processors.reset` / `parts.reset_processor` clears it."""

    def __init__(self, num_action_repeats=1, additional_discount=0.99):
        self._repeats, self._gamma = num_action_repeats, additional_discount
        self.reset()

    def reset(self):
        self._window = collections.deque(maxlen=STACK)
        self._count, self._reward, self._discount = 0, 0.0, 1.0

    def __call__(self, timestep):
        if timestep.first():
            self.reset()
            self._window.append(timestep.observation)
            return TimeStep(0, None, None, stack_frames(self._window))
        self._count += 1
        self._reward += timestep.reward
        self._discount *= timestep.discount
        if self._count < self._repeats and not timestep.last():
            return None
        self._window.append(timestep.observation)
        out = TimeStep(timestep.step_type, self._reward, self._discount * self._gamma,
                       stack_frames(self._window))
        self._count, self._reward, self._discount = 0, 0.0, 1.0
        return out
