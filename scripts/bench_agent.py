"""C1 stand-in (ALE is absent and outside the optimisation target): the DoubleDqn drop-in agent driven by
`parts.run_loop` on the synthetic environment with the reference's schedule (action repeat 4, learn
every 16 frames, batch 32, FIFO replay).  Reports environment frames/s and the cost of the acting
call (`q_values`: one batch-1 forward + host epsilon-greedy) on its own.
   python scripts/bench_agent.py [frames] [replay capacity] [precision]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import json
import numpy as np
from dqn_mgsc_zoo_b200 import agent as agent_lib, networks, parts, replay as rp, synthetic

FRAMES = int(sys.argv[1]) if len(sys.argv) > 1 else 40_000
CAP = int(sys.argv[2]) if len(sys.argv) > 2 else 100_000
PREC = sys.argv[3] if len(sys.argv) > 3 else "tf32"
A = 6
replay = rp.TransitionReplay(CAP, rp.Transition(None, None, None, None, None), np.random.RandomState(0))
ag = agent_lib.DoubleDqn(
    preprocessor=synthetic.StackingPreprocessor(num_action_repeats=4),
    sample_network_input=np.zeros((84, 84, 4), np.uint8), network=networks.double_dqn_atari_network(A),
    optimizer=networks.rmsprop(learning_rate=0.00025, decay=0.95, eps=0.01 / 32 ** 2, centered=True),
    transition_accumulator=rp.TransitionAccumulator(), replay=replay, batch_size=32,
    exploration_epsilon=lambda t: 0.1, min_replay_capacity_fraction=0.02, learn_period=16,
    target_network_update_period=30_000, grad_error_bound=1.0 / 32, rng_key=1, precision=PREC)
env = synthetic.SyntheticEnvironment(np.random.default_rng(3), A, min_len=50, max_len=400)

# environment + preprocessing alone (no agent): the host floor of the loop
class _Null(parts.Agent):
    def __init__(self): self.pre = synthetic.StackingPreprocessor(num_action_repeats=4)
    def step(self, ts): self.pre(ts); return 0
    def reset(self): self.pre.reset()
    def get_state(self): return {}
    def set_state(self, s): pass
    @property
    def statistics(self): return {}
t0 = time.perf_counter()
for i, _ in enumerate(parts.run_loop(_Null(), env, max_steps_per_episode=27_000)):
    if i == FRAMES // 4: break
env_fps = (FRAMES // 4) / (time.perf_counter() - t0)

warm = int(0.02 * CAP) * 4 + 2000
n, t0, learn0 = 0, None, 0
for _ in parts.run_loop(ag, env, max_steps_per_episode=27_000):
    n += 1
    if n == warm:
        ag.learner.synchronize()
        t0 = time.perf_counter()
    if n == warm + FRAMES:
        break
ag.learner.synchronize()
dt = time.perf_counter() - t0
obs = np.random.default_rng(0).integers(0, 256, (84, 84, 4), dtype=np.uint8)
for _ in range(50): ag.learner.q_values(obs)
t1 = time.perf_counter()
for _ in range(2000): ag.learner.q_values(obs)
act_us = (time.perf_counter() - t1) / 2000 * 1e6
print(json.dumps({"agent_frames_per_s": FRAMES / dt, "frames": FRAMES, "learner_updates_per_s": FRAMES / 16 / dt,
                  "env_plus_preprocessor_only_frames_per_s": env_fps, "q_values_call_us": act_us,
                  "replay_capacity": CAP, "precision": PREC,
                  "schedule": "action repeat 4, learn every 16 frames, batch 32, epsilon 0.1"}))
