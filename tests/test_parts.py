"""CPU tests of the host seam (dqn_mgsc_zoo_b200/parts.py): the run-loop call-order tape of
the reference's parts_test.py:80-166 and its LinearSchedule checks (:37-75), restated on the
dm_env-free TimeStep stand-in."""
import numpy as np
import pytest

from dqn_mgsc_zoo_b200 import parts
from dqn_mgsc_zoo_b200 import synthetic


class DummyAgent(parts.Agent):
    """test_utils.py:30-58."""

    def __init__(self, tape):
        self._tape = tape

    def reset(self):
        self._tape.append('Agent reset')

    def step(self, timestep):
        self._tape.append('Agent step')
        return 0

    def get_state(self):
        return {}

    def set_state(self, state):
        pass

    @property
    def statistics(self):
        return {}


class DummyEnvironment:
    """test_utils.py:61-96."""

    def __init__(self, tape, episode_length):
        self._tape, self._episode_length = tape, episode_length

    def reset(self):
        self._t = 0
        self._tape.append('Environment reset')
        return synthetic.TimeStep(parts.FIRST, 0.0, 0.0, 1.0)

    def step(self, action):
        self._tape.append('Environment step (%s)' % action)
        self._t += 1
        if self._t == self._episode_length:
            step_type, self._t = parts.LAST, -1
        else:
            step_type = parts.FIRST if self._t == 0 else parts.MID
        return synthetic.TimeStep(step_type, 2.0, 0.0 if step_type == parts.LAST else 1.0, 1.0)


def test_run_loop_tape_matches_reference():
    tape = []
    agent, env = DummyAgent(tape), DummyEnvironment(tape, episode_length=4)
    episode_index, t, max_steps = 0, 0, 14
    for _, timestep_t, _, _ in parts.run_loop(agent, env, max_steps_per_episode=100, yield_before_reset=True):
        tape.append((episode_index, t, timestep_t is None))
        if timestep_t is None:
            tape.append('Episode begin')
            continue
        if timestep_t.last():
            tape.append('Episode end')
            episode_index += 1
        if t + 1 >= max_steps:
            tape.append('Maximum number of steps reached')
            break
        t += 1

    def episode(e, t0, steps):
        out = [(e, t0, True), 'Episode begin', 'Agent reset', 'Environment reset', 'Agent step', (e, t0, False)]
        for k in range(1, steps):
            out += ['Environment step (0)', 'Agent step', (e, t0 + k, False)]
        return out

    expected = (episode(0, 0, 5) + ['Episode end'] + episode(1, 5, 5) + ['Episode end'] + episode(2, 10, 4)
                + ['Maximum number of steps reached'])
    assert tape == expected


def test_run_loop_truncates_long_episodes():
    tape = []
    agent, env = DummyAgent(tape), DummyEnvironment(tape, episode_length=50)
    seen = []
    for _, ts, _, a in parts.run_loop(agent, env, max_steps_per_episode=3):
        seen.append((ts.step_type, a))
        if len(seen) == 8:
            break
    # steps 0..2 with actions, then the truncated LAST timestep with a_t None; then a new episode
    assert seen[:4] == [(parts.FIRST, 0), (parts.MID, 0), (parts.MID, 0), (parts.LAST, None)]
    assert seen[4] == (parts.FIRST, 0)


def test_linear_schedule():
    s = parts.LinearSchedule(begin_t=5, end_t=12, begin_value=1.0, end_value=0.3)
    for step in range(20):
        want = 1.0 if step <= 5 else (0.3 if step >= 12 else 1.0 - (step - 5) / 7 * 0.7)
        assert abs(s(step) - want) < 1e-12
    s2 = parts.LinearSchedule(begin_t=5, decay_steps=7, begin_value=0.5, end_value=0.5)
    assert all(abs(s2(t) - 0.5) < 1e-12 for t in range(20))
    with pytest.raises(ValueError, match='Exactly one of'):
        parts.LinearSchedule(begin_value=0.0, end_value=1.0, begin_t=5)
    with pytest.raises(ValueError, match='Exactly one of'):
        parts.LinearSchedule(begin_value=0.0, end_value=1.0, begin_t=5, end_t=12, decay_steps=7)


def test_epsilon_greedy_distribution():
    rng = np.random.default_rng(0)
    q = np.array([0.1, 0.9, 0.3, 0.9])           # two tied maxima share the greedy mass
    n = 40000
    counts = np.bincount([parts.epsilon_greedy(rng, q, 0.2) for _ in range(n)], minlength=4) / n
    want = np.array([0.05, 0.05 + 0.4, 0.05, 0.05 + 0.4])
    assert np.abs(counts - want).max() < 0.01


def test_stacking_preprocessor_layout():
    """A000, AB00, ABC0, ABCD, BCDE (processors.py:489-505) and action repeat."""
    pre = synthetic.StackingPreprocessor(num_action_repeats=1)
    frames = [np.full((84, 84), k + 1, np.uint8) for k in range(6)]
    out = [pre(synthetic.TimeStep(0, None, None, frames[0]))]
    for f in frames[1:]:
        out.append(pre(synthetic.TimeStep(1, 1.0, 1.0, f)))
    chans = [tuple(int(o.observation[0, 0, c]) for c in range(4)) for o in out]
    assert chans == [(1, 0, 0, 0), (1, 2, 0, 0), (1, 2, 3, 0), (1, 2, 3, 4), (2, 3, 4, 5), (3, 4, 5, 6)]
    pre4 = synthetic.StackingPreprocessor(num_action_repeats=4)
    assert pre4(synthetic.TimeStep(0, None, None, frames[0])) is not None
    got = [pre4(synthetic.TimeStep(1, 1.0, 1.0, f)) for f in frames[1:5]]
    assert [g is None for g in got] == [True, True, True, False]
    assert got[3].reward == 4.0 and abs(got[3].discount - 0.99) < 1e-12
