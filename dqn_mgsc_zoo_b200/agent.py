"""Drop-in DQN / Double-DQN agents whose `_learn` runs on the B200 learner.

Mirrors dqn/agent.py:40-229 (`Dqn`) and double_q/agent.py:40-233 (`DoubleDqn`): same
constructor keywords, `step` / `reset` / `get_state` / `set_state`, `statistics`,
`online_params`, `exploration_epsilon`, same learn / target-sync schedule (dqn/agent.py:149-156),
same error on a repeat before any action.  What changes is below the seam:

  * `network` is a `networks.NatureCnnSpec` and `optimizer` a `networks.RmsPropSpec` (the
    haiku / optax objects of the reference cannot exist without those packages);
  * `_learn` (dqn/agent.py:179-189) hands the sampled replay *slots* to
    `dqnb_learner_update_from_replay`: gather+stack, forward x2/x3, TD, backward, RMSProp run
    as device work with the batch never leaving HBM.  A replay that is not backed by the HBM
    frame store still works: its host `sample()` batch goes through `dqnb_learner_update_host`;
  * the target sync (`:155-156`) is a device copy;
  * acting (`:121-131,169-177`) runs the batch-1 forward on the device and draws the
    eps-greedy action from a NumPy Generator seeded from `rng_key` (same distribution as
    distrax.EpsilonGreedy; not the same threefry stream).
"""
import collections
from typing import Any, Callable, Mapping

import numpy as np

from dqn_mgsc_zoo_b200 import learner as lrn
from dqn_mgsc_zoo_b200 import networks
from dqn_mgsc_zoo_b200 import parts

# optax.rmsprop(centered=True) state is (ScaleByRStdDevState(mu, nu), ScaleByScheduleState/EmptyState);
# get_state() keeps that shape with haiku-shaped trees of NumPy arrays inside.
ScaleByRStdDevState = collections.namedtuple("ScaleByRStdDevState", "mu nu")
EmptyState = collections.namedtuple("EmptyState", "")


class Dqn(parts.Agent):
    """Deep Q-Network agent (dqn/agent.py:40)."""

    _DOUBLE_Q = False
    _ROW_WEIGHTED = 0

    def __init__(
        self,
        preprocessor,
        sample_network_input,
        network: networks.NatureCnnSpec,
        optimizer: networks.RmsPropSpec,
        transition_accumulator: Any,
        replay,
        batch_size: int,
        exploration_epsilon: Callable[[int], float],
        min_replay_capacity_fraction: float,
        learn_period: int,
        target_network_update_period: int,
        grad_error_bound: float,
        rng_key,
        device: int = 0,
        precision: str = "tf32",
    ):
        """`device`, `precision`: this implementation's two extra keywords.  precision="tf32" (default) runs every
        conv / dense GEMM on the tcgen05 tensor cores (fp32 storage, tf32 operands, fp32 accumulate; parity class
        rel 1e-3, the mode bench.py reports); precision="fp32" is the CUDA-core path held to rel 1e-5, 3.5x slower."""
        if not optimizer.centered:
            raise ValueError("the CUDA learner implements optax.rmsprop(centered=True) only "
                             "(dqn/run_atari.py:208-213)")
        if tuple(np.shape(sample_network_input)) != (84, 84, 4):
            raise ValueError("the CUDA learner is built for 84x84x4 observations, got %s"
                             % (np.shape(sample_network_input),))
        self._preprocessor = preprocessor
        self._replay = replay
        self._transition_accumulator = transition_accumulator
        self._batch_size = batch_size
        self._exploration_epsilon = exploration_epsilon
        self._min_replay_capacity = min_replay_capacity_fraction * replay.capacity
        self._learn_period = learn_period
        self._target_network_update_period = target_network_update_period
        self._network = network
        self.precision = precision

        # jax.random.split(rng_key) -> (agent key, network init key) in the reference.
        seeds = np.random.SeedSequence(networks.key_to_seed(rng_key)).spawn(2)
        self._rng = np.random.default_rng(seeds[0])
        self._learner = lrn.Learner(
            network.num_actions, batch_size, shared_bias=network.shared_bias, double_q=self._DOUBLE_Q,
            learning_rate=optimizer.learning_rate, decay=optimizer.decay, eps=optimizer.eps,
            grad_error_bound=grad_error_bound, precision=precision, device=device,
            row_weighted=self._ROW_WEIGHTED)
        init = networks.init_params(np.random.default_rng(seeds[1]), network.num_actions, network.shared_bias)
        self._learner.set_params(init, lrn.ONLINE)
        self._learner.sync_target()           # self._target_params = self._online_params

        self._action = None
        self._frame_t = -1
        self._statistics = {'state_value': np.nan}

    # ------------------------------------------------------------------ parts.Agent
    def step(self, timestep) -> parts.Action:
        """Selects action given timestep and potentially learns (dqn/agent.py:133-158)."""
        self._frame_t += 1

        timestep = self._preprocessor(timestep)

        if timestep is None:  # Repeat action.
            if self._action is None:
                raise RuntimeError('Cannot repeat if action has never been selected.')
            action = self._action
        else:
            action = self._action = self._act(timestep)
            for transition in self._transition_accumulator.step(timestep, action):
                self._replay.add(transition)

        if self._replay.size < self._min_replay_capacity:
            return action

        if self._frame_t % self._learn_period == 0:
            self._learn()

        if self._frame_t % self._target_network_update_period == 0:
            self._learner.sync_target()

        return action

    def reset(self) -> None:
        self._transition_accumulator.reset()
        parts.reset_processor(self._preprocessor)
        self._action = None

    def _act(self, timestep) -> parts.Action:
        q_t = self._learner.q_values(timestep.observation)[0]
        self._statistics['state_value'] = float(q_t.max())
        return parts.Action(parts.epsilon_greedy(self._rng, q_t, self.exploration_epsilon))

    def _learn(self) -> None:
        store = getattr(self._replay, "frame_store", None)
        if store is not None and hasattr(self._replay, "sample_slots"):
            self._learner.update_from_replay(store, self._replay.sample_slots(self._batch_size))
        else:
            t = self._replay.sample(self._batch_size)
            self._learner.update_host(t.s_tm1, t.a_tm1, t.r_t, t.discount_t, t.s_t)

    @property
    def online_params(self):
        """Current Q-network parameters as a haiku-shaped tree of NumPy arrays (D2H copy)."""
        return networks.to_haiku_tree(self._learner.get_params(lrn.ONLINE))

    @property
    def statistics(self) -> Mapping[str, float]:
        return self._statistics

    @property
    def exploration_epsilon(self) -> float:
        return self._exploration_epsilon(self._frame_t)

    @property
    def learner(self) -> lrn.Learner:
        return self._learner

    def get_state(self) -> Mapping[str, Any]:
        """Same keys as dqn/agent.py:210-220; device arrays come back as NumPy."""
        tree = lambda which: networks.to_haiku_tree(self._learner.get_params(which))
        return {
            'rng_key': self._rng.bit_generator.state,
            'frame_t': self._frame_t,
            'opt_state': (ScaleByRStdDevState(mu=tree(lrn.MU), nu=tree(lrn.NU)), EmptyState()),
            'online_params': tree(lrn.ONLINE),
            'target_params': tree(lrn.TARGET),
            'replay': self._replay.get_state(),
        }

    def set_state(self, state: Mapping[str, Any]) -> None:
        self._rng.bit_generator.state = state['rng_key']
        self._frame_t = state['frame_t']
        rms = state['opt_state'][0]
        self._learner.set_params(networks.from_haiku_tree(rms.mu), lrn.MU)
        self._learner.set_params(networks.from_haiku_tree(rms.nu), lrn.NU)
        self._learner.set_params(networks.from_haiku_tree(state['online_params']), lrn.ONLINE)
        self._learner.set_params(networks.from_haiku_tree(state['target_params']), lrn.TARGET)
        self._replay.set_state(state['replay'])


class DoubleDqn(Dqn):
    """Double DQN (tuned) agent (double_q/agent.py:40): the online net on s_t picks the action,
    the target net values it (rlax.double_q_learning, double_q/agent.py:85-111)."""

    _DOUBLE_Q = True


class PrioritizedDqn(Dqn):
    """Prioritized Experience Replay (proportional variant) DQN agent (prioritized/agent.py:40-262).

    Same constructor keywords as the reference; `replay` is a `replay.PrioritizedTransitionReplay`.  Differences
    from `DoubleDqn` follow the reference line by line: new transitions enter with the largest priority seen so
    far (:147), `_learn` (:215-236) samples (transitions, ids, importance weights), runs the double-Q update with
    loss = mean(w * 0.5 td^2) - the weight multiplies the cotangent before rlax.clip_gradient (:85-108;
    `row_weighted = 2` in the learner) - and writes |td| back as the new priorities.  The td errors come back to
    the host every step (32 floats), as in the reference (`jax.device_get(priorities)`, :230): the next
    sample depends on them, so this agent is host-synchronous by construction.
    """

    _DOUBLE_Q = True
    _ROW_WEIGHTED = 2

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self._max_seen_priority = 1.0

    def step(self, timestep) -> parts.Action:
        self._frame_t += 1

        timestep = self._preprocessor(timestep)

        if timestep is None:  # Repeat action.
            if self._action is None:
                raise RuntimeError('Cannot repeat if action has never been selected.')
            action = self._action
        else:
            action = self._action = self._act(timestep)
            for transition in self._transition_accumulator.step(timestep, action):
                self._replay.add(transition, priority=self._max_seen_priority)

        if self._replay.size < self._min_replay_capacity:
            return action

        if self._frame_t % self._learn_period == 0:
            self._learn()

        if self._frame_t % self._target_network_update_period == 0:
            self._learner.sync_target()

        return action

    def _learn(self) -> None:
        store = getattr(self._replay, "frame_store", None)
        if store is not None and hasattr(self._replay, "sample_slots_ids_weights"):
            slots, indices, weights = self._replay.sample_slots_ids_weights(self._batch_size)
            self._learner.update_from_replay(store, slots, weights=weights)
        else:
            t, indices, weights = self._replay.sample(self._batch_size)
            self._learner.set_row_weights(weights)
            self._learner.update_host(t.s_tm1, t.a_tm1, t.r_t, t.discount_t, t.s_t)
        priorities = np.abs(self._learner.read(lrn.TD))
        max_priority = priorities.max()
        self._max_seen_priority = np.max([self._max_seen_priority, max_priority])
        self._replay.update_priorities(indices, priorities)

    @property
    def importance_sampling_exponent(self) -> float:
        return self._replay.importance_sampling_exponent

    @property
    def max_seen_priority(self) -> float:
        return self._max_seen_priority

    def get_state(self) -> Mapping[str, Any]:
        state = dict(super().get_state())
        state['max_seen_priority'] = self._max_seen_priority
        return state

    def set_state(self, state: Mapping[str, Any]) -> None:
        super().set_state(state)
        self._max_seen_priority = state['max_seen_priority']


ScaleByAdamState = collections.namedtuple("ScaleByAdamState", "count mu nu")


class MGSCDqn(Dqn):
    """MGSC agent (dqn_mgsc_batched/agent.py:40-398): DQN whose replay sampling logits are
    meta-learned.  Extra constructor keywords `meta_optimizer` (networks.AdamSpec, the reference
    passes optax.adam, dqn_mgsc_batched/run_atari.py:241-243) and `meta_batch_size`.

    Per step (agent.py:252-278): meta-learn on the newest transition *before* it is added to the
    replay, then add, learn from a logit-weighted sample, sync.  `meta_update` (agent.py:160-220)
    runs on the device (dqnb_mgsc_meta_update: one row-weighted batched backward, two virtual RMSProp steps,
    softmax backward, Adam on the logits); neither params nor RMSProp state advance in it.

    VARIANT: `second_order=False` (default) is the STOP-GRADIENT meta objective of dqn_mgsc_batched
    (agent.py:189-191, target `stop_gradient(theta_exp + rmsprop(g'))`).  `second_order=True` is the objective of the
    sibling dqn_mgsc_batched_reservoir/agent.py:160-198, which omits that stop_gradient: the meta loss is
    differentiated through the gradient at the expected parameters (a Hessian-vector product through the CNN,
    dqnb_learner_hvp) and through the optimiser state; `MGSCReservoirDqn` below is that agent.  The attribute
    `meta_variant` ('stop_gradient' / 'second_order') says which one runs (`statistics` must hold floats for the
    reference's trackers).
    """

    _SECOND_ORDER = False

    def __init__(self, *args, meta_optimizer: networks.AdamSpec, meta_batch_size: int, second_order=None, **kwargs):
        super().__init__(*args, **kwargs)
        self._meta_batch_size = meta_batch_size
        second_order = self._SECOND_ORDER if second_order is None else bool(second_order)
        self._meta = lrn.MetaLearner(self._learner, meta_batch_size, learning_rate=meta_optimizer.learning_rate,
                                     b1=meta_optimizer.b1, b2=meta_optimizer.b2, eps=meta_optimizer.eps,
                                     second_order=second_order)
        self._last_transitions = []
        self._statistics['meta_loss'] = np.nan
        self.meta_variant = 'second_order' if second_order else 'stop_gradient'      # see the class docstring

    def step(self, timestep) -> parts.Action:
        self._frame_t += 1

        timestep = self._preprocessor(timestep)

        transitions = []
        if timestep is None:  # Repeat action.
            if self._action is None:
                raise RuntimeError('Cannot repeat if action has never been selected.')
            action = self._action
        else:
            action = self._action = self._act(timestep)
            transitions = [t for t in self._transition_accumulator.step(timestep, action)]
            self._last_transitions = self._last_transitions + transitions

        # Meta-prioritization learn step (agent.py:252-262)
        if (self._frame_t % self._learn_period == 0) and (
                self._replay.size >= max(self._meta_batch_size, self._min_replay_capacity)):
            if len(self._last_transitions) != 0:
                self._meta_prioritization_learn(self._last_transitions[-1])
                self._last_transitions.clear()
        if timestep is not None:
            for transition in transitions:
                self._replay.add(transition)

        if self._replay.size < self._min_replay_capacity:
            return action

        if self._frame_t % self._learn_period == 0:
            self._learn()

        if self._frame_t % self._target_network_update_period == 0:
            self._learner.sync_target()

        return action

    def reset(self) -> None:
        super().reset()
        self._last_transitions = []

    def _meta_prioritization_learn(self, online_transition) -> None:
        """agent.py:307-345: draw a uniform meta batch, meta_update, write the logits back."""
        store = getattr(self._replay, "frame_store", None)
        if store is not None and hasattr(self._replay, "batch_of_ids_slots_and_logits"):
            indices, slots, logits = self._replay.batch_of_ids_slots_and_logits(self._meta_batch_size)
            new_logits, loss = self._meta.meta_update_from_replay(store, slots, logits, online_transition)
        else:
            indices, transitions, logits = self._replay.batch_of_ids_transitions_and_logits(self._meta_batch_size)
            new_logits, loss = self._meta.meta_update(transitions, logits, online_transition)
        self._statistics['meta_loss'] = loss
        self._replay.update_priorities(indices, new_logits)

    @property
    def meta_learner(self) -> lrn.MetaLearner:
        return self._meta

    def get_state(self) -> Mapping[str, Any]:
        state = dict(super().get_state())
        m = self._meta.get_state()
        state['meta_opt_state'] = (ScaleByAdamState(count=m['count'], mu=m['mu'], nu=m['nu']), EmptyState())
        return state

    def set_state(self, state: Mapping[str, Any]) -> None:
        super().set_state(state)
        adam = state['meta_opt_state'][0]
        self._meta.set_state({'count': adam.count, 'mu': adam.mu, 'nu': adam.nu})


class MGSCReservoirDqn(MGSCDqn):
    """The MGSCDqn of dqn_mgsc_batched_reservoir/agent.py:42-402: same agent, meta loss WITHOUT stop_gradient on the
    target parameters (:183-195), meant for `replay_circular.MGSCReservoirTransitionReplay`
    (dqn_mgsc_batched_reservoir/run_atari.py:211)."""

    _SECOND_ORDER = True
