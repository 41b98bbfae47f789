"""Host-side logic of the data-parallel learner (dqn_mgsc_zoo_b200/parallel.py) on CPU:
world_size-2 gloo processes, the oracle standing in for the CUDA learner (tests may use it),
checked against the single-process oracle update at the global batch."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from dqn_mgsc_zoo_b200 import networks
from dqn_mgsc_zoo_b200 import parallel
from dqn_mgsc_zoo_b200 import synthetic
from oracle import learner_oracle as net
from oracle import update_oracle as upd

A, GLOBAL_B, STEPS = 4, 8, 2


class OracleShardLearner:
    """The learner interface DataParallelLearner drives, computed by the float64 oracle."""

    def __init__(self, params, batch_size, global_batch, double_q):
        self.online = {k: v.astype(np.float64) for k, v in params.items()}
        self.target = {k: v.copy() for k, v in self.online.items()}
        self.batch_size, self.global_batch, self.double_q = batch_size, global_batch, double_q
        P = net.flatten(self.online).size
        self.mu, self.nu = np.zeros(P), np.zeros(P)
        self.grad = torch.zeros(P, dtype=torch.float64)

    def update_host(self, s0, a, r, d, s1):
        assert len(a) == self.batch_size
        out = upd.loss_and_grads(self.online, self.target, (s0, a, r, d, s1), self.double_q,
                                 mean_over=self.global_batch)
        self.grad.copy_(torch.from_numpy(net.flatten(out["grads"])))

    def apply_grads(self):
        A_, shared = self.online["fc2.w"].shape[1], self.online["fc2.b"].shape == (1,)
        p, self.mu, self.nu = upd.rmsprop_step(net.flatten(self.online), self.grad.numpy().copy(), self.mu, self.nu)
        self.online = net.unflatten(p, A_, shared)


def _batches():
    rng = np.random.default_rng(3)
    return [synthetic.random_batch(rng, GLOBAL_B, A) for _ in range(STEPS)]


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        params = networks.init_params(np.random.default_rng(0), A, True)
        lo, hi = parallel.shard_bounds(GLOBAL_B, world, rank)
        local = OracleShardLearner(params, hi - lo, GLOBAL_B, double_q=True)
        dp = parallel.DataParallelLearner(local, GLOBAL_B, rank, world, local.grad,
                                          lambda t: dist.all_reduce(t, op=dist.ReduceOp.SUM))
        for batch in _batches():
            dp.update_host(*batch)
        np.save(os.path.join(out_dir, "rank%d.npy" % rank), net.flatten(local.online))
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_update_equals_global_batch_update(tmp_path):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    params = networks.init_params(np.random.default_rng(0), A, True)
    online = {k: v.astype(np.float64) for k, v in params.items()}
    target = {k: v.copy() for k, v in online.items()}
    P = net.flatten(online).size
    mu, nu = np.zeros(P), np.zeros(P)
    for batch in _batches():
        out = upd.update(online, target, mu, nu, batch, double_q=True)
        online, mu, nu = out["params"], out["mu"], out["nu"]
    want = net.flatten(online)
    r0, r1 = np.load(tmp_path / "rank0.npy"), np.load(tmp_path / "rank1.npy")
    assert np.array_equal(r0, r1)                                  # replicas stay identical
    step = np.abs(want - net.flatten({k: v.astype(np.float64) for k, v in params.items()})).max()
    assert np.abs(r0 - want).max() < 1e-9 * step + 1e-15           # float64: only summation order differs


def test_shard_bounds_and_slots():
    assert parallel.shard_bounds(4096, 8, 3) == (1536, 2048)
    with pytest.raises(ValueError, match="does not split"):
        parallel.shard_bounds(10, 4, 0)
    with pytest.raises(ValueError, match="not in a world"):
        parallel.shard_bounds(8, 2, 2)

    class Stub:
        batch_size = 4
    dp = parallel.DataParallelLearner(Stub(), 8, 1, 2, None, None)
    assert np.array_equal(dp.local_slots(np.arange(8) * 10), [40, 50, 60, 70])
    with pytest.raises(ValueError, match="need 8 global slots"):
        dp.local_slots(np.arange(7))
    with pytest.raises(ValueError, match="local learner batch"):
        parallel.DataParallelLearner(Stub(), 16, 0, 2, None, None)
    arr = parallel.DeviceArray(0x1000, 7)
    assert arr.__cuda_array_interface__["shape"] == (7,)
