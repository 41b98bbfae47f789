"""Steady-state period of update_from_replay (batch 32, tf32) against the gather's CTA split, with the
gather overlapped on its own stream: how much does the prefetching gather cost the update it runs beside?"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import bench
from dqn_mgsc_zoo_b200 import learner as lrn, networks, replay as rp, _lib
cap = 60000
replay = rp.TransitionReplay(cap, rp.Transition(None, None, None, None, None), np.random.RandomState(0))
src = bench.EpisodeSource(5)
for _ in range(cap):
    replay.add(rp.Transition(*src.next()))
store = replay.frame_store
store.flush()
L = lrn.Learner(6, 32, shared_bias=True, double_q=True, precision="tf32")
init = networks.init_params(np.random.default_rng(0), 6, True)
L.set_params(init, lrn.ONLINE); L.set_params(init, lrn.TARGET)
slots = [replay.sample_slots(32) for _ in range(64)]
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for splits in (0, 16, 8, 5, 3, 2, 1):
    _lib.call("dqnb_replay_set_gather_splits", store._h, splits)
    for i in range(100): L.update_from_replay(store, slots[i % 64])
    torch.cuda.synchronize(); e0.record()
    for i in range(2000): L.update_from_replay(store, slots[i % 64])
    e1.record(); torch.cuda.synchronize()
    print("gather splits %2d (0 = default): %.2f us/update" % (splits, e0.elapsed_time(e1) / 2000 * 1e3))
