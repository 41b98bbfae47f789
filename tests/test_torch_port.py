"""The timed CPU baseline (oracle/torch_port.py) computes the same update as the oracle."""
import numpy as np

from dqn_mgsc_zoo_b200 import synthetic
from oracle import learner_oracle as net
from oracle import torch_port
from oracle import update_oracle as upd


def test_torch_port_matches_oracle():
    rng = np.random.default_rng(0)
    params = net.init_params(rng, 6, True)
    port = torch_port.TorchCpuLearner(params, double_q=True)
    batch = synthetic.random_batch(rng, 8, 6)
    P = net.flatten(params).size
    want = upd.update(params, params, np.zeros(P), np.zeros(P), batch, double_q=True)
    loss = port.update(*batch)
    assert abs(loss - want["loss"]) < 1e-5 * max(1.0, abs(want["loss"]))
    got = port.params_np()
    for k, v in want["params"].items():
        step_ref = v - params[k]
        step = got[k].astype(np.float64) - params[k]
        scale = np.abs(step_ref).max()
        assert np.abs(step - step_ref).max() < 2e-3 * scale + 2e-7, k
