"""GPU tests of the deferred-apply seam and the NCCL data-parallel learner.

  * 1 GPU: update(defer_apply) + apply_grads == the fused update, gradient arena == oracle's;
  * >= 2 GPUs (skipped otherwise; `gpurun --gpus 2`): two ranks over NCCL on half batches with the
    global loss mean == one learner on the whole batch (rel 1e-5: only summation order differs),
    replicas bit-identical to each other.
"""
import os
import socket

import numpy as np
import pytest

from dqn_mgsc_zoo_b200 import learner as lrn
from dqn_mgsc_zoo_b200 import networks
from dqn_mgsc_zoo_b200 import parallel
from dqn_mgsc_zoo_b200 import synthetic

pytestmark = pytest.mark.gpu
A, STEPS = 6, 2


def rel(x, ref):
    x, ref = np.asarray(x, dtype=np.float64), np.asarray(ref, dtype=np.float64)
    return np.abs(x - ref).max() / (np.abs(ref).max() + 1e-30)


@pytest.mark.parametrize("precision", ["fp32", "tf32"])
def test_deferred_apply_equals_fused_update(precision):
    B = 32
    init = networks.init_params(np.random.default_rng(0), A, True)
    batches = [synthetic.random_batch(np.random.default_rng(10 + i), B, A) for i in range(3)]
    fused = lrn.Learner(A, B, shared_bias=True, double_q=True, precision=precision)
    split = lrn.Learner(A, B, shared_bias=True, double_q=True, precision=precision, defer_apply=True)
    for L in (fused, split):
        L.set_params(init, lrn.ONLINE)
        L.set_params(init, lrn.TARGET)
    for b in batches:
        before = split.read(lrn.ONLINE)
        fused.update_host(*b)
        split.update_host(*b)
        assert np.array_equal(split.read(lrn.ONLINE), before)      # nothing applied yet
        g_split = split.read(lrn.GRAD)
        assert rel(g_split, fused.read(lrn.GRAD)) < 1e-6
        split.apply_grads()
        # same gradient values, same optimiser arithmetic (fp32: bit-identical kernels; tf32: the fused
        # path applies fc1 inside its GEMM epilogue, same formula)
        for which in (lrn.ONLINE, lrn.MU, lrn.NU):
            assert rel(split.read(which), fused.read(which)) < 1e-6, which
    ptr, n = split.device_ptr(lrn.GRAD)
    assert ptr and n == split.num_params


TAPS = (110, 111, 112, 113)


def _worker(rank, world, port, out_dir, precision, global_b, in_library):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        dp = parallel.make_nccl_learner(A, global_b, rank, rank, world, shared_bias=True, double_q=True,
                                        precision=precision, in_library=in_library)
        assert dp.in_graph == in_library
        L = dp.learner
        init = networks.init_params(np.random.default_rng(0), A, True)
        L.set_params(init, lrn.ONLINE)
        L.set_params(init, lrn.TARGET)
        out = {}
        for i in range(STEPS):
            dp.update_host(*synthetic.random_batch(np.random.default_rng(100 + i), global_b, A))
            torch.cuda.synchronize()
            if i == 0:      # what the oracle comparison needs: the GLOBAL gradient and this rank's forward decisions
                out["grad"], out["td"], out["online1"] = L.read(lrn.GRAD), L.read(lrn.TD), L.read(lrn.ONLINE)
                for tap in TAPS:
                    out["tap%d" % tap] = L.read(tap)
        out["online"] = L.read(lrn.ONLINE)
        out["desc"] = np.array(dp.describe())
        np.savez(os.path.join(out_dir, "rank%d.npz" % rank), **out)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("precision,global_b,in_library", [("fp32", 64, False), ("fp32", 64, True),
                                                           ("tf32", 512, True), ("tf32", 512, False)])
def test_two_gpu_nccl_update_matches_oracle(tmp_path, precision, global_b, in_library):
    """Two ranks over NCCL, half the rows each, global loss mean: (1) the all-reduced gradient and the parameter
    step equal the ORACLE's update on the whole batch (fp32: rel 1e-5 raw; tf32 at 256 rows per rank = the
    persistent stream GEMM kernels: rel 1e-3 through the ranks' own ReLU masks / td errors, tests/cond.py);
    (2) the replicas stay bit-identical; (3) in_library: the collective runs inside the update graph
    (dqnb_dp_init), else host-dispatched torch.distributed.all_reduce."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    import torch.multiprocessing as mp
    from oracle import learner_oracle as net
    from oracle import update_oracle as upd
    from tests import cond
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(2, port, str(tmp_path), precision, global_b, in_library), nprocs=2, join=True)
    r = [np.load(tmp_path / ("rank%d.npz" % k)) for k in range(2)]
    assert np.array_equal(r[0]["online"], r[1]["online"]) and np.array_equal(r[0]["grad"], r[1]["grad"])
    assert ("inside the update graph" in str(r[0]["desc"])) == in_library
    init = networks.init_params(np.random.default_rng(0), A, True)
    f64 = {k: v.astype(np.float64) for k, v in init.items()}
    batch = synthetic.random_batch(np.random.default_rng(100), global_b, A)
    P = r[0]["grad"].size
    g_dev = r[0]["grad"].astype(np.float64)
    if precision == "fp32":
        want = upd.update(f64, f64, np.zeros(P), np.zeros(P), batch, True)
        assert rel(g_dev, want["flat_grad"]) < 1e-5
        assert rel(r[0]["online1"].astype(np.float64) - net.flatten(f64), net.flatten(want["params"]) - net.flatten(f64)) < 1e-4
    else:
        read = lambda tap: np.concatenate([r[0]["tap%d" % tap], r[1]["tap%d" % tap]])
        td = np.concatenate([r[0]["td"], r[1]["td"]]).astype(np.float64)
        g_ref = cond.grads(f64, cond.device_cache(read, batch[0], global_b), td, batch[1], global_b)
        assert rel(g_dev, g_ref) < 1e-3
        p2, _, _ = upd.rmsprop_step(net.flatten(f64), g_dev, np.zeros(P), np.zeros(P))
        assert rel(r[0]["online1"].astype(np.float64) - net.flatten(f64), p2 - net.flatten(f64)) < 2e-3
        q_ref = net.forward(f64, batch[0][:4])
        assert np.isfinite(q_ref).all()
