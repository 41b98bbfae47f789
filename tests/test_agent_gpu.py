"""GPU test of the drop-in agents (dqn_mgsc_zoo_b200/agent.py) driven through
`parts.run_loop`, in lockstep with a CPU oracle agent that follows dqn/agent.py:133-189 with
the oracle replay (oracle/replay_oracle.py) and the float64 oracle `update`."""
import pickle

import numpy as np
import pytest

from dqn_mgsc_zoo_b200 import agent as agent_lib
from dqn_mgsc_zoo_b200 import learner as lrn
from dqn_mgsc_zoo_b200 import networks
from dqn_mgsc_zoo_b200 import parts
from dqn_mgsc_zoo_b200 import replay as rp
from dqn_mgsc_zoo_b200 import synthetic
from oracle import learner_oracle as net
from oracle import replay_oracle as ro
from oracle import update_oracle as upd

pytestmark = pytest.mark.gpu
A, B, CAP = 6, 8, 64


def rel(x, ref):
    x, ref = np.asarray(x, dtype=np.float64), np.asarray(ref, dtype=np.float64)
    return np.abs(x - ref).max() / (np.abs(ref).max() + 1e-30)


def make_agent(cls, replay, seed=7, eps=1.0, reservoir=False, precision="fp32"):
    # fp32 (SIMT) is the parity mode of these lock-step tests; the agents themselves default to tf32
    spec = networks.double_dqn_atari_network(A) if cls._DOUBLE_Q else networks.dqn_atari_network(A)
    return cls(
        preprocessor=synthetic.StackingPreprocessor(num_action_repeats=2),
        sample_network_input=np.zeros((84, 84, 4), np.uint8), network=spec,
        optimizer=networks.rmsprop(learning_rate=0.00025, decay=0.95, eps=0.01 / 32 ** 2, centered=True),
        transition_accumulator=rp.TransitionAccumulator(), replay=replay, batch_size=B,
        exploration_epsilon=lambda t: eps, min_replay_capacity_fraction=0.25, learn_period=4,
        target_network_update_period=24, grad_error_bound=1.0 / 32, rng_key=seed, precision=precision)


class OracleAgent:
    """dqn/agent.py:133-189 on the CPU oracle; takes its actions from a tape (the CUDA agent's)."""

    def __init__(self, online, replay, double_q):
        self.pre = synthetic.StackingPreprocessor(num_action_repeats=2)
        self.acc = ro.PairAccumulator()
        self.replay, self.double_q = replay, double_q
        self.online = {k: v.astype(np.float64) for k, v in online.items()}
        self.target = {k: v.copy() for k, v in self.online.items()}
        P = net.flatten(self.online).size
        self.mu, self.nu = np.zeros(P), np.zeros(P)
        self.frame_t, self.batches, self.syncs = -1, [], 0

    def reset(self):
        self.acc.reset()
        self.pre.reset()

    def step(self, timestep, action):
        self.frame_t += 1
        ts = self.pre(timestep)
        if ts is not None:
            for tr in self.acc.step(ts, action):
                self.replay.add(tr)
        if self.replay.size < 0.25 * CAP:
            return
        if self.frame_t % 4 == 0:
            batch = self.replay.sample(B)
            self.batches.append(batch)
            out = upd.update(self.online, self.target, self.mu, self.nu,
                             (batch.s_tm1, batch.a_tm1, batch.r_t, batch.discount_t, batch.s_t), self.double_q)
            self.online, self.mu, self.nu = out["params"], out["mu"], out["nu"]
        if self.frame_t % 24 == 0:
            self.target = {k: v.copy() for k, v in self.online.items()}
            self.syncs += 1


@pytest.mark.parametrize("cls,reservoir", [(agent_lib.Dqn, False), (agent_lib.DoubleDqn, False),
                                           (agent_lib.Dqn, True)])
def test_agent_run_loop_tracks_oracle(cls, reservoir):
    structure = rp.Transition(None, None, None, None, None)
    ours_replay = (rp.ReservoirTransitionReplay if reservoir else rp.TransitionReplay)(
        CAP, structure, np.random.RandomState(3))
    ref_replay = (ro.ReservoirReplay if reservoir else ro.FifoReplay)(
        CAP, ro.Transition(None, None, None, None, None), np.random.RandomState(3))
    ag = make_agent(cls, ours_replay)
    ora = OracleAgent(ag.learner.get_params(lrn.ONLINE), ref_replay, cls is agent_lib.DoubleDqn)
    # record every batch the CUDA agent learns from (slots -> gathered host arrays)
    seen = []
    real = ours_replay.sample_slots

    def spy(n):
        slots = real(n)
        seen.append(ours_replay.frame_store.gather_host(slots))
        return slots
    ours_replay.sample_slots = spy

    env = synthetic.SyntheticEnvironment(np.random.default_rng(11), A, min_len=3, max_len=40)
    frames = 0
    for _, timestep, _, action in parts.run_loop(ag, env, max_steps_per_episode=30, yield_before_reset=True):
        if timestep is None:
            ora.reset()
            continue
        # the oracle agent sees the same timestep and the action the CUDA agent chose/repeated
        ora.step(timestep, ag._action)
        frames += 1
        if frames >= 260:
            break
    assert len(seen) == len(ora.batches) >= 30 and ora.syncs >= 5
    for got, want in zip(seen, ora.batches):     # bit-exact batches, same random stream
        assert np.array_equal(got[0], want.s_tm1) and np.array_equal(got[4], want.s_t)
        assert np.array_equal(got[1], want.a_tm1)
        assert np.array_equal(got[2], np.asarray(want.r_t, np.float32))
        assert np.array_equal(got[3], np.asarray(want.discount_t, np.float32))
    assert ours_replay.size == ref_replay.size and ag._frame_t == ora.frame_t
    # free-running fp32 vs float64 over 30+ updates (see test_chained_updates_free_running)
    got = ag.learner.get_params(lrn.ONLINE)
    for k, v in ora.online.items():
        assert rel(got[k], v) < 5e-2, k
    assert rel(ag.learner.read(lrn.TARGET), net.flatten(ora.target)) < 5e-2
    assert np.isfinite(ag.statistics['state_value'])


def test_agent_errors_and_properties():
    replay = rp.TransitionReplay(CAP, rp.Transition(None, None, None, None, None), np.random.RandomState(0))
    ag = make_agent(agent_lib.Dqn, replay, eps=0.0)
    ag.reset()
    pre = ag._preprocessor
    # a skipped (None) timestep before any action was selected
    first = synthetic.TimeStep(0, None, None, np.zeros((84, 84), np.uint8))
    mid = synthetic.TimeStep(1, 0.0, 1.0, np.zeros((84, 84), np.uint8))
    ag._preprocessor = lambda ts: None
    with pytest.raises(RuntimeError, match='Cannot repeat if action has never been selected.'):
        ag.step(mid)
    ag._preprocessor = pre
    a = ag.step(first)
    assert isinstance(a, int) and 0 <= a < A
    # greedy action == argmax of the oracle forward on the same params
    q = net.forward({k: v.astype(np.float64) for k, v in ag.learner.get_params().items()},
                    synthetic.stack_frames([first.observation])[None])
    assert a == int(np.argmax(q[0])) and abs(ag.statistics['state_value'] - q.max()) < 1e-5
    tree = ag.online_params
    assert set(tree) == set(networks.NATURE_HAIKU_NAMES.values())
    assert tree['sequential/sequential/conv2_d']['w'].shape == (8, 8, 4, 32)
    assert ag.exploration_epsilon == 0.0


def test_agent_state_round_trip_resumes_identically():
    def fresh():
        replay = rp.TransitionReplay(CAP, rp.Transition(None, None, None, None, None), np.random.RandomState(5))
        return make_agent(agent_lib.DoubleDqn, replay, seed=9, eps=0.3), replay

    def drive(ag, env, n):
        acts = []
        for _, ts, _, a in parts.run_loop(ag, env, max_steps_per_episode=25):
            acts.append(a)
            if len(acts) >= n:
                return acts

    ag1, rep1 = fresh()
    env_rng = np.random.default_rng(21)
    env = synthetic.SyntheticEnvironment(env_rng, A, min_len=3, max_len=40)
    drive(ag1, env, 120)
    state = pickle.loads(pickle.dumps({"agent": ag1.get_state(), "rs": rep1._random_state.get_state(),
                                       "env": env_rng.bit_generator.state}))
    assert set(state["agent"]) == {'rng_key', 'frame_t', 'opt_state', 'online_params', 'target_params', 'replay'}
    ag2, rep2 = fresh()
    ag2.set_state(state["agent"])
    rep2._random_state.set_state(state["rs"])
    env_rng2 = np.random.default_rng(0)
    env_rng2.bit_generator.state = state["env"]
    env2 = synthetic.SyntheticEnvironment(env_rng2, A, min_len=3, max_len=40)
    a1 = drive(ag1, env, 80)
    a2 = drive(ag2, env2, 80)
    assert a1 == a2
    assert np.array_equal(ag1.learner.read(lrn.ONLINE), ag2.learner.read(lrn.ONLINE))
    assert np.array_equal(ag1.learner.read(lrn.NU), ag2.learner.read(lrn.NU))


def test_epsilon_greedy_actor_uses_online_params():
    replay = rp.TransitionReplay(CAP, rp.Transition(None, None, None, None, None), np.random.RandomState(0))
    ag = make_agent(agent_lib.Dqn, replay)
    actor = parts.EpsilonGreedyActor(synthetic.StackingPreprocessor(), networks.dqn_atari_network(A), 0.0, rng_key=1)
    actor.network_params = ag.online_params
    actor.reset()
    obs = np.random.default_rng(2).integers(0, 256, (84, 84), dtype=np.uint8)
    a = actor.step(synthetic.TimeStep(0, None, None, obs))
    q = ag.learner.q_values(synthetic.stack_frames([obs]))[0]
    assert a == int(np.argmax(q))
    st = actor.get_state()
    assert set(st) == {'rng_key', 'network_params'}
    actor.set_state(pickle.loads(pickle.dumps(st)))



def test_agents_default_to_the_tensor_core_learner():
    replay = rp.TransitionReplay(CAP, rp.Transition(None, None, None, None, None), np.random.RandomState(0))
    ag = agent_lib.DoubleDqn(
        preprocessor=synthetic.StackingPreprocessor(num_action_repeats=2),
        sample_network_input=np.zeros((84, 84, 4), np.uint8), network=networks.double_dqn_atari_network(A),
        optimizer=networks.rmsprop(learning_rate=0.00025, decay=0.95, eps=0.01 / 32 ** 2, centered=True),
        transition_accumulator=rp.TransitionAccumulator(), replay=replay, batch_size=B,
        exploration_epsilon=lambda t: 1.0, min_replay_capacity_fraction=0.25, learn_period=4,
        target_network_update_period=24, grad_error_bound=1.0 / 32, rng_key=3)
    assert ag.precision == "tf32"


class OraclePrioritizedAgent:
    """prioritized/agent.py:133-236 on the CPU oracle; actions come from the CUDA agent's tape."""

    def __init__(self, online, replay):
        self.pre = synthetic.StackingPreprocessor(num_action_repeats=2)
        self.acc = ro.PairAccumulator()
        self.replay = replay
        self.online = {k: v.astype(np.float64) for k, v in online.items()}
        self.target = {k: v.copy() for k, v in self.online.items()}
        P = net.flatten(self.online).size
        self.mu, self.nu = np.zeros(P), np.zeros(P)
        self.frame_t, self.max_seen, self.log = -1, 1.0, []

    def reset(self):
        self.acc.reset()
        self.pre.reset()

    def step(self, timestep, action, device_td=None):
        self.frame_t += 1
        ts = self.pre(timestep)
        if ts is not None:
            for tr in self.acc.step(ts, action):
                self.replay.add(tr, self.max_seen)
        if self.replay.size < 0.25 * CAP:
            return
        if self.frame_t % 4 == 0:
            batch, ids, w = self.replay.sample(B)
            out = upd.update(self.online, self.target, self.mu, self.nu,
                             (batch.s_tm1, batch.a_tm1, batch.r_t, batch.discount_t, batch.s_t), True, weights=w)
            self.log.append((ids, w, out["td"], out["loss"], net.flatten(self.online), self.mu, self.nu, batch))
            self.online, self.mu, self.nu = out["params"], out["mu"], out["nu"]
            # priorities: |td| as float32, exactly what the device hands back (teacher-forced with the device's
            # own td so that both replays keep sampling the same ids; the td values themselves are compared)
            pri = np.abs(np.asarray(device_td if device_td is not None else out["td"], dtype=np.float32))
            self.max_seen = np.max([self.max_seen, pri.max()])
            self.replay.update_priorities(ids, pri)
        if self.frame_t % 24 == 0:
            self.target = {k: v.copy() for k, v in self.online.items()}


def test_prioritized_agent_tracks_oracle():
    """PrioritizedDqn in lock-step with an oracle agent over the oracle prioritized replay: sampled ids
    bit-exact, importance weights rel 1e-6 (float64 on both sides: exact in practice), td errors and the
    importance-weighted loss at the fp32 tolerance; priorities are teacher-forced with the device's |td|."""
    beta = lambda t: min(1.0, 0.4 + 0.002 * t)
    structure = rp.Transition(None, None, None, None, None)
    ours = rp.PrioritizedTransitionReplay(CAP, structure, 0.6, beta, 0.1, True, np.random.RandomState(5))
    ref = ro.PrioritizedReplay(CAP, ro.Transition(None, None, None, None, None), 0.6, beta, 0.1, True,
                               np.random.RandomState(5))
    ag = make_agent(agent_lib.PrioritizedDqn, ours)
    ora = OraclePrioritizedAgent(ag.learner.get_params(lrn.ONLINE), ref)
    seen = []
    real_sample, real_update = ours.sample_slots_ids_weights, ours.update_priorities

    def spy_sample(n):
        slots, ids, w = real_sample(n)
        state = (ag.learner.read(lrn.ONLINE), ag.learner.read(lrn.MU), ag.learner.read(lrn.NU),
                 ag.learner.read(lrn.TARGET))
        seen.append({"ids": np.array(ids), "w": np.array(w), "batch": ours.frame_store.gather_host(slots),
                     "state": state})
        return slots, ids, w

    def spy_update(ids, pri):
        seen[-1]["pri"] = np.array(pri)
        seen[-1]["loss"] = ag.learner.loss()
        real_update(ids, pri)
    ours.sample_slots_ids_weights, ours.update_priorities = spy_sample, spy_update

    env = synthetic.SyntheticEnvironment(np.random.default_rng(13), A, min_len=3, max_len=40)
    frames = 0
    for _, timestep, _, action in parts.run_loop(ag, env, max_steps_per_episode=30, yield_before_reset=True):
        if timestep is None:
            ora.reset()
            continue
        n_before = len(ora.log)
        # the device learnt during ag.step (already done inside run_loop); hand its |td| to the oracle replay
        ora.step(timestep, ag._action, device_td=seen[len(ora.log)]["pri"] if len(seen) > len(ora.log) else None)
        assert len(ora.log) - n_before in (0, 1)
        frames += 1
        if frames >= 240:
            break
    assert len(seen) == len(ora.log) >= 25
    for k, (got, want) in enumerate(zip(seen, ora.log)):
        ids, w, td, loss, theta, mu, nu, batch = want
        assert np.array_equal(got["ids"], ids), k
        assert rel(got["w"], w) < 1e-6, k
        assert np.array_equal(got["batch"][0], batch.s_tm1) and np.array_equal(got["batch"][4], batch.s_t)
        # one-step parity from the device's own state (teacher-forced like test_chained_updates_teacher_forced)
        on = net.unflatten(got["state"][0].astype(np.float64), A, True)
        tgt = net.unflatten(got["state"][3].astype(np.float64), A, True)
        out = upd.update(on, tgt, got["state"][1].astype(np.float64), got["state"][2].astype(np.float64),
                         (batch.s_tm1, batch.a_tm1, batch.r_t, batch.discount_t, batch.s_t), True, weights=got["w"])
        assert rel(got["pri"], np.abs(out["td"])) < 1e-5, k
        assert abs(got["loss"] - out["loss"]) < 1e-5 * max(1.0, abs(out["loss"])), k
    assert ag.max_seen_priority >= 1.0 and 0.4 <= ag.importance_sampling_exponent <= 1.0
    assert ours.check_valid() == (True, '')
    state = ag.get_state()
    assert state['max_seen_priority'] == ag.max_seen_priority and 'distribution' in state['replay']
