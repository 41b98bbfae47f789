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
A, GLOBAL_B, STEPS = 6, 64, 3


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


def _worker(rank, world, port, out_dir):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        dp = parallel.make_nccl_learner(A, GLOBAL_B, rank, rank, world, shared_bias=True, double_q=True)
        init = networks.init_params(np.random.default_rng(0), A, True)
        dp.learner.set_params(init, lrn.ONLINE)
        dp.learner.set_params(init, lrn.TARGET)
        for i in range(STEPS):
            dp.update_host(*synthetic.random_batch(np.random.default_rng(100 + i), GLOBAL_B, A))
        torch.cuda.synchronize()
        np.save(os.path.join(out_dir, "rank%d.npy" % rank), dp.learner.read(lrn.ONLINE))
    finally:
        dist.destroy_process_group()


def test_two_gpu_nccl_update_equals_global_batch_update(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    single = lrn.Learner(A, GLOBAL_B, shared_bias=True, double_q=True)
    init = networks.init_params(np.random.default_rng(0), A, True)
    single.set_params(init, lrn.ONLINE)
    single.set_params(init, lrn.TARGET)
    start = single.read(lrn.ONLINE)
    for i in range(STEPS):
        single.update_host(*synthetic.random_batch(np.random.default_rng(100 + i), GLOBAL_B, A))
    want = single.read(lrn.ONLINE)
    r0, r1 = np.load(tmp_path / "rank0.npy"), np.load(tmp_path / "rank1.npy")
    assert np.array_equal(r0, r1)
    assert rel(r0 - start, want - start) < 1e-4       # parameter *step*; params themselves agree to ~1e-8
    assert rel(r0, want) < 1e-6
