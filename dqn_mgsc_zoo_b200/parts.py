"""Host-side seam of the learner hot path: the pieces of dqn_zoo/parts.py an agent and its
caller need so that the CUDA learner is a drop-in.

Mirrors (same names, arguments, call order, errors):
  Agent               parts.py:43-68     abstract interface (step/reset/get_state/set_state/statistics)
  run_loop            parts.py:71-123    agent <-> environment alternation (tape pinned by parts_test.py:80-166)
  EpsilonGreedyActor  parts.py:343-412   evaluation actor on externally set network params
  LinearSchedule      parts.py:415-431   exploration epsilon schedule

Trackers, CSV writers and the pickle Checkpoint of parts.py are host bookkeeping outside the
path and are not rebuilt; the reference's own work unchanged on top of
these agents because they only use the interface above.
"""
import abc
from typing import Any, Iterable, Mapping, Optional, Tuple

import numpy as np

Action = int

FIRST, MID, LAST = 0, 1, 2   # dm_env.StepType values


def _as_last(timestep):
    """timestep._replace(step_type=LAST) for dm_env.TimeStep (enum) or the int stand-in."""
    step_type = timestep.step_type
    last = getattr(type(step_type), "LAST", LAST)
    return timestep._replace(step_type=last)


class Agent(abc.ABC):
    """Agent interface (parts.py:43-68)."""

    @abc.abstractmethod
    def step(self, timestep) -> Action:
        """Selects action given timestep and potentially learns."""

    @abc.abstractmethod
    def reset(self) -> None:
        """Resets the agent's episodic state such as frame stack and action repeat."""

    @abc.abstractmethod
    def get_state(self) -> Mapping[str, Any]:
        """Retrieves agent state as a dictionary (e.g. for serialization)."""

    @abc.abstractmethod
    def set_state(self, state: Mapping[str, Any]) -> None:
        """Sets agent state from a (potentially de-serialized) dictionary."""

    @property
    @abc.abstractmethod
    def statistics(self) -> Mapping[str, float]:
        """Returns current agent statistics as a dictionary."""


def run_loop(agent: Agent, environment, max_steps_per_episode: int = 0,
             yield_before_reset: bool = False) -> Iterable[Tuple[Any, Optional[Any], Agent, Optional[Action]]]:
    """Repeatedly alternates step calls on environment and agent (parts.py:71-123).

    Yields `(environment, timestep_t, agent, a_t)` with `a_t = agent.step(timestep_t)`; on the
    LAST timestep of an episode (or when `max_steps_per_episode` truncates it) the agent gets
    one extra step whose action is dropped, and `a_t` is None.
    """
    while True:
        if yield_before_reset:
            yield environment, None, agent, None

        t = 0
        agent.reset()
        timestep_t = environment.reset()

        while True:
            a_t = agent.step(timestep_t)
            yield environment, timestep_t, agent, a_t

            t += 1
            timestep_t = environment.step(a_t)

            if max_steps_per_episode > 0 and t >= max_steps_per_episode:
                assert t == max_steps_per_episode
                timestep_t = _as_last(timestep_t)

            if timestep_t.last():
                agent.step(timestep_t)   # extra agent step, action ignored
                yield environment, timestep_t, agent, None
                break


def reset_processor(processor) -> None:
    """processors.reset (processors.py:48-51): calls `reset()` if the processor has one."""
    if hasattr(processor, "reset"):
        processor.reset()


def epsilon_greedy(rng: np.random.Generator, q_values: np.ndarray, epsilon: float) -> int:
    """One draw from distrax.EpsilonGreedy(q, eps) (dqn/agent.py:121-129): probability
    eps/A on every action plus (1-eps) shared equally by the argmax ties.  The reference
    samples with a threefry key; here a NumPy Generator, so actions agree in distribution,
    not draw-for-draw."""
    q = np.asarray(q_values)
    if rng.random() < epsilon:
        return int(rng.integers(0, q.shape[-1]))
    best = np.flatnonzero(q == q.max())
    return int(best[0] if best.size == 1 else rng.choice(best))


class EpsilonGreedyActor(Agent):
    """Acts eps-greedily with network params set from outside (parts.py:343-412).

    `network` is a `networks.NatureCnnSpec`; the forward pass runs on the CUDA learner's
    batch-1 path (dqnb_learner_q_values).  `network_params` takes the haiku-shaped tree that
    `Dqn.online_params` returns.
    """

    def __init__(self, preprocessor, network, exploration_epsilon: float, rng_key, device: int = 0):
        from dqn_mgsc_zoo_b200 import learner as lrn
        from dqn_mgsc_zoo_b200 import networks
        self._preprocessor = preprocessor
        self._network = network
        self._exploration_epsilon = exploration_epsilon
        self._rng = np.random.default_rng(networks.key_to_seed(rng_key))
        self._action = None
        self._forward = lrn.Learner(network.num_actions, 1, shared_bias=network.shared_bias, device=device)
        self._network_params = None

    @property
    def network_params(self):
        return self._network_params

    @network_params.setter
    def network_params(self, params):
        from dqn_mgsc_zoo_b200 import learner as lrn
        from dqn_mgsc_zoo_b200 import networks
        self._network_params = params
        if params is not None:
            self._forward.set_params(networks.from_haiku_tree(params), lrn.ONLINE)

    def step(self, timestep) -> Action:
        timestep = self._preprocessor(timestep)
        if timestep is None:
            if self._action is None:
                raise RuntimeError('Cannot repeat if action has never been selected.')
            return self._action
        q_t = self._forward.q_values(timestep.observation)[0]
        self._action = Action(epsilon_greedy(self._rng, q_t, self._exploration_epsilon))
        return self._action

    def reset(self) -> None:
        reset_processor(self._preprocessor)
        self._action = None

    def get_state(self) -> Mapping[str, Any]:
        return {'rng_key': self._rng.bit_generator.state, 'network_params': self._network_params}

    def set_state(self, state: Mapping[str, Any]) -> None:
        self._rng.bit_generator.state = state['rng_key']
        self.network_params = state['network_params']

    @property
    def statistics(self) -> Mapping[str, float]:
        return {}


class LinearSchedule:
    """Linear schedule, used for exploration epsilon in DQN agents (parts.py:415-431)."""

    def __init__(self, begin_value, end_value, begin_t, end_t=None, decay_steps=None):
        if (end_t is None) == (decay_steps is None):
            raise ValueError('Exactly one of end_t, decay_steps must be provided.')
        self._decay_steps = decay_steps if end_t is None else end_t - begin_t
        self._begin_t = begin_t
        self._begin_value = begin_value
        self._end_value = end_value

    def __call__(self, t):
        frac = min(max(t - self._begin_t, 0), self._decay_steps) / self._decay_steps
        return (1 - frac) * self._begin_value + frac * self._end_value
