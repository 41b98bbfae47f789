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
  * an all-reduce (sum, fp32, 1 687 206 elements = 6.75 MB) over the flat gradient arena.  On the GPU box it
    lives INSIDE the library (`dqnb_dp_init`): the learner's CUDA graph sums the split-K partials of
    [fc1.w, end) - 95 % of the bytes, ready first in backward order - and all-reduces them with NCCL on a side
    stream while the convolution backward still runs, then the convolution gradients, then applies the step:
    one graph launch per rank, no Python between backward and optimiser.  The host-dispatched variant
    (`torch.distributed.all_reduce` between `update` and `apply_grads`) remains for the gloo CPU tests and as
    the comparison arm (`in_library=False`);
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

    def __init__(self, learner, global_batch: int, rank: int, world: int, grad_buffer, all_reduce_sum,
                 in_graph: bool = False):
        self.learner, self.global_batch, self.rank, self.world = learner, int(global_batch), rank, world
        self.in_graph = bool(in_graph)     # the learner's own graph all-reduces and applies (dqnb_dp_init)
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
        if self.in_graph:
            return
        if self.world > 1:
            self._all_reduce_sum(self.grad_buffer)
        self.learner.apply_grads()

    def all_reduce_only(self):
        """The collective(s) of one update alone (timing aid; sums whatever the gradient arena holds)."""
        if self.in_graph:
            self.learner.dp_allreduce()
        elif self.world > 1:
            self._all_reduce_sum(self.grad_buffer)

    def describe(self) -> str:
        if self.world == 1:
            return "single rank: no collective"
        if self.in_graph:
            return ("ncclAllReduce x2 (fc1+fc2 chunk, then conv chunk) captured inside the update graph on a "
                    "high-priority side stream, NCCL %s" % self.learner.dp_info()[2])
        return "host-dispatched torch.distributed.all_reduce between update and apply_grads"

    def close(self):
        self.learner.close()


def make_nccl_learner(num_actions, global_batch, device, rank, world, group=None, in_library=True,
                      **learner_kwargs):
    """CUDA learner whose gradient arena is all-reduced over NCCL.

    in_library=True (default): `dqnb_dp_init` - the collective is part of the learner's CUDA graph; torch.distributed
    is used once, to hand rank 0's ncclUniqueId to the other ranks.  in_library=False: the round-1 arrangement - the
    learner runs on the legacy default stream, which is torch's current stream, so ProcessGroupNCCL orders a
    host-dispatched all_reduce after the backward kernels and the optimiser step after it.
    """
    import torch
    import torch.distributed as dist
    from dqn_mgsc_zoo_b200 import _lib
    from dqn_mgsc_zoo_b200 import learner as lrn

    lo, hi = shard_bounds(global_batch, world, rank)
    L = lrn.Learner(num_actions, hi - lo, loss_mean_over=global_batch, defer_apply=True, device=device,
                    **learner_kwargs)
    if in_library and world > 1:
        uid = np.zeros(128, dtype=np.uint8)
        if rank == 0:
            _lib.call("dqnb_dp_unique_id", uid.ctypes.data)
        t = torch.from_numpy(uid).to(torch.device("cuda", device))
        dist.broadcast(t, src=0, group=group)
        L.dp_init(t.cpu().numpy(), rank, world)
        return DataParallelLearner(L, global_batch, rank, world, None, None, in_graph=True)
    ptr, n = L.device_ptr(lrn.GRAD)
    grad = torch.as_tensor(DeviceArray(ptr, n), device=torch.device("cuda", device))
    assert grad.data_ptr() == ptr and grad.numel() == n

    def all_reduce_sum(t):
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)

    return DataParallelLearner(L, global_batch, rank, world, grad, all_reduce_sum)
