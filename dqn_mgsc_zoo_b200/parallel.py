"""Large-batch data-parallel learner: the one exchange step of the path.

The reference has no multi-device code; this is how its `update` (dqn/agent.py:109-117) is spread
over GPUs without changing its arithmetic:

  * the global index vector of B_global transitions is drawn once (same RandomState stream on every
    rank, i.e. exactly the draw a single learner with batch B_global would make); rank r takes
    slots [r*B_local, (r+1)*B_local) and gathers them from its own replica of the replay ring
    (adds are replicated - 7 KB per step - so gathers stay local);
  * every rank runs forward / TD / backward on its rows with the *global* loss mean
    (`loss_mean_over = B_global`: rlax.clip_gradient clips td/B_global, dqn/agent.py:94-106), so
    the local flat gradient is this rank's share of the global sum;
  * ONE all-reduce (sum, fp32, 1 687 206 elements = 6.75 MB) over the flat gradient arena -
    NCCL over NVLink 5 / NVSwitch on the GPU box, gloo in the CPU tests;
  * every rank applies the same centered-RMSProp step to its replica of params / mu / nu
    (`dqnb_learner_apply_grads`), so the replicas stay bit-identical to each other.

fp32 summation order differs from a single GPU, so parity with the B_global single-device update is
rel 1e-5, not bit-exact; rank-to-rank state is bit-identical (same reduced gradient everywhere).
"""
import numpy as np


def shard_bounds(global_batch: int, world: int, rank: int):
    """Rows [lo, hi) of the global batch that `rank` owns; the batch must split evenly."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("rank %d not in a world of %d" % (rank, world))
    if global_batch % world != 0:
        raise ValueError("global batch %d does not split over %d ranks" % (global_batch, world))
    local = global_batch // world
    return rank * local, (rank + 1) * local


class DeviceArray:
    """A raw device pointer dressed up for torch.as_tensor (CUDA array interface v2)."""

    def __init__(self, ptr: int, n: int, typestr: str = "<f4"):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 2}


class DataParallelLearner:
    """Wraps a local learner created with `defer_apply=True, loss_mean_over=global_batch`.

    `learner` needs update_from_replay / update_host / apply_grads; `grad_buffer` is whatever
    `all_reduce_sum` accepts in place (a CUDA tensor aliasing the gradient arena for NCCL, a CPU
    tensor in the gloo tests).
    """

    def __init__(self, learner, global_batch: int, rank: int, world: int, grad_buffer, all_reduce_sum):
        self.learner, self.global_batch, self.rank, self.world = learner, int(global_batch), rank, world
        self.lo, self.hi = shard_bounds(self.global_batch, world, rank)
        if getattr(learner, "batch_size", self.hi - self.lo) != self.hi - self.lo:
            raise ValueError("local learner batch %d != shard size %d" % (learner.batch_size, self.hi - self.lo))
        self.grad_buffer, self._all_reduce_sum = grad_buffer, all_reduce_sum

    def local_slots(self, global_slots):
        s = np.asarray(global_slots)
        if s.shape[0] != self.global_batch:
            raise ValueError("need %d global slots, got %d" % (self.global_batch, s.shape[0]))
        return s[self.lo:self.hi]

    def update_from_replay(self, frame_store, global_slots):
        self.learner.update_from_replay(frame_store, self.local_slots(global_slots))
        self._finish()

    def update_host(self, s_tm1, a_tm1, r_t, discount_t, s_t):
        """Takes the GLOBAL batch (every rank holds it, like the replicated replay) and uses its rows."""
        sl = slice(self.lo, self.hi)
        self.learner.update_host(s_tm1[sl], a_tm1[sl], r_t[sl], discount_t[sl], s_t[sl])
        self._finish()

    def _finish(self):
        if self.world > 1:
            self._all_reduce_sum(self.grad_buffer)
        self.learner.apply_grads()


def make_nccl_learner(num_actions, global_batch, device, rank, world, group=None, **learner_kwargs):
    """CUDA learner + torch.distributed (NCCL) all-reduce on its gradient arena.

    The learner runs on the legacy default stream, which is torch's current stream, so
    ProcessGroupNCCL orders the collective after the backward kernels and the optimiser step
    after the collective without host synchronisation.
    """
    import torch
    import torch.distributed as dist
    from dqn_mgsc_zoo_b200 import learner as lrn

    lo, hi = shard_bounds(global_batch, world, rank)
    L = lrn.Learner(num_actions, hi - lo, loss_mean_over=global_batch, defer_apply=True, device=device,
                    **learner_kwargs)
    ptr, n = L.device_ptr(lrn.GRAD)
    grad = torch.as_tensor(DeviceArray(ptr, n), device=torch.device("cuda", device))
    assert grad.data_ptr() == ptr and grad.numel() == n

    def all_reduce_sum(t):
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)

    return DataParallelLearner(L, global_batch, rank, world, grad, all_reduce_sum)
