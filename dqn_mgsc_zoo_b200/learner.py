"""Python handle over the C-ABI learner (dqnb_learner_*): the device-side `update`."""
import collections
import ctypes

import numpy as np

from dqn_mgsc_zoo_b200 import _lib

ONLINE, TARGET, MU, NU, GRAD, Q_TM1, Q_TARGET_T, Q_T, TD, LOSS, ROW_WEIGHT = range(11)
PRECISIONS = {"fp32": 0, "tf32": 2}


class LearnerConfig(ctypes.Structure):
    _fields_ = [("num_actions", ctypes.c_int32), ("shared_bias", ctypes.c_int32),
                ("double_q", ctypes.c_int32), ("batch_size", ctypes.c_int32),
                ("loss_mean_over", ctypes.c_int32), ("precision", ctypes.c_int32),
                ("learning_rate", ctypes.c_float), ("decay", ctypes.c_float),
                ("eps", ctypes.c_float), ("grad_error_bound", ctypes.c_float),
                ("defer_apply", ctypes.c_int32), ("row_weighted", ctypes.c_int32)]


def param_shapes(num_actions, shared_bias=False):
    """Leaf shapes in arena order (haiku names in networks.NATURE_HAIKU_NAMES)."""
    s = collections.OrderedDict()
    s["conv1.w"], s["conv1.b"] = (8, 8, 4, 32), (32,)
    s["conv2.w"], s["conv2.b"] = (4, 4, 32, 64), (64,)
    s["conv3.w"], s["conv3.b"] = (3, 3, 64, 64), (64,)
    s["fc1.w"], s["fc1.b"] = (3136, 512), (512,)
    s["fc2.w"], s["fc2.b"] = (512, num_actions), ((1,) if shared_bias else (num_actions,))
    return s


def flatten_params(params):
    return np.concatenate([np.asarray(v, dtype=np.float32).ravel() for v in params.values()])


def unflatten_params(flat, num_actions, shared_bias=False):
    out, o = collections.OrderedDict(), 0
    for k, shape in param_shapes(num_actions, shared_bias).items():
        n = int(np.prod(shape))
        out[k] = np.asarray(flat[o:o + n]).reshape(shape)
        o += n
    return out


class Learner:
    def __init__(self, num_actions, batch_size, shared_bias=False, double_q=False, learning_rate=0.00025,
                 decay=0.95, eps=0.01 / 32 ** 2, grad_error_bound=1.0 / 32, precision="fp32", device=0,
                 stream=0, loss_mean_over=0, defer_apply=False, row_weighted=False):
        self.num_actions, self.batch_size = int(num_actions), int(batch_size)
        self.shared_bias, self.double_q = bool(shared_bias), bool(double_q)
        self.device, self.stream = device, stream
        cfg = LearnerConfig(self.num_actions, int(self.shared_bias), int(self.double_q), self.batch_size,
                            int(loss_mean_over), PRECISIONS[precision], learning_rate, decay, eps,
                            grad_error_bound, int(bool(defer_apply)), int(row_weighted))
        self.defer_apply = bool(defer_apply)
        h = ctypes.c_void_p()
        _lib.call("dqnb_learner_create", ctypes.byref(cfg), device, ctypes.byref(h))
        self._h = h
        self._loss_outstanding = 0
        n = ctypes.c_int64()
        _lib.call("dqnb_learner_num_params", self._h, ctypes.byref(n))
        self.num_params = n.value
        self._update_from_replay = _lib.load().dqnb_learner_update_from_replay
        # per-step calls: bound once, their out-parameter allocated once (a fresh c_float + byref + name lookup per
        # call were a third of loss_lagged's host time)
        fc = _lib.fastcall()
        self._fc = fc
        if fc is not None:
            self._fc_update = (_lib.fn_address("dqnb_learner_update_from_replay"), self._h.value)
            self._fc_loss = (_lib.fn_address("dqnb_learner_loss_post"), _lib.fn_address("dqnb_learner_loss_take"),
                             self._h.value)
        self._loss_post_fn = _lib.load().dqnb_learner_loss_post
        self._loss_take_fn = _lib.load().dqnb_learner_loss_take
        self._loss_out = ctypes.c_float()
        self._loss_out_ref = ctypes.byref(self._loss_out)

    def close(self):
        if self._h is not None:
            _lib.call("dqnb_learner_destroy", self._h)
            self._h = None
            self._fc = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _len(self, which):
        if which <= GRAD:
            return self.num_params
        if which in (Q_TM1, Q_TARGET_T, Q_T):
            return self.batch_size * self.num_actions
        dbg = {100: 512, 101: 3136, 102: 5184, 103: 12800, 110: 512, 111: 3136, 112: 5184, 113: 12800}
        if which in dbg:
            return self.batch_size * dbg[which]
        return self.batch_size if which == TD else 1

    def write(self, which, array):
        a = np.ascontiguousarray(array, dtype=np.float32).ravel()
        _lib.call("dqnb_learner_write", self._h, which, a.ctypes.data, a.size, self.stream)

    def read(self, which):
        out = np.empty(self._len(which), dtype=np.float32)
        _lib.call("dqnb_learner_read", self._h, which, out.ctypes.data, out.size, self.stream)
        if which in (Q_TM1, Q_TARGET_T, Q_T):
            return out.reshape(self.batch_size, self.num_actions)
        return out

    def set_params(self, params, which=ONLINE):
        self.write(which, flatten_params(params))

    def get_params(self, which=ONLINE):
        return unflatten_params(self.read(which), self.num_actions, self.shared_bias)

    def loss(self):
        return float(self.read(LOSS)[0])

    def device_ptr(self, which):
        """(device address, length) of a state array - e.g. the flat gradient arena for NCCL."""
        ptr, n = ctypes.c_void_p(), ctypes.c_int64()
        _lib.call("dqnb_learner_device_ptr", self._h, which, ctypes.byref(ptr), ctypes.byref(n))
        return ptr.value, n.value

    def apply_grads(self):
        """optimizer.update + apply_updates on the gradient arena (second half of a deferred update)."""
        _lib.call("dqnb_learner_apply_grads", self._h, self.stream)

    def dp_init(self, unique_id, rank, world):
        """Joins the NCCL communicator described by rank 0's 128-byte ncclUniqueId; afterwards every update of
        this (defer_apply) learner all-reduces the gradient arena inside its own CUDA graph and applies it."""
        uid = np.ascontiguousarray(unique_id, dtype=np.uint8)
        if uid.size != 128:
            raise ValueError("ncclUniqueId is 128 bytes, got %d" % uid.size)
        _lib.call("dqnb_dp_init", self._h, uid.ctypes.data, int(rank), int(world))

    def dp_allreduce(self):
        _lib.call("dqnb_dp_allreduce", self._h, self.stream)

    def dp_info(self):
        """(rank, world, 'major.minor.patch' of the NCCL in use)"""
        r, w, v = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        _lib.call("dqnb_dp_info", self._h, ctypes.byref(r), ctypes.byref(w), ctypes.byref(v))
        ver = v.value
        return r.value, w.value, "%d.%d.%d" % (ver // 10000, (ver // 100) % 100, ver % 100)

    def loss_post(self):
        """Queues a read-back of the latest update's loss behind it (does not wait)."""
        _lib.check(self._loss_post_fn(self._h, self.stream))
        self._loss_outstanding += 1

    def loss_take(self):
        """Waits for the oldest queued loss read-back and returns it."""
        _lib.check(self._loss_take_fn(self._h, self._loss_out_ref))
        self._loss_outstanding -= 1
        return self._loss_out.value

    def loss_lagged(self):
        """loss_post() for this update, then the previous update's loss (None the first time):
        per-step statistics without ever idling the device."""
        if self._fc is not None:
            post, take, h = self._fc_loss
            want = self._loss_outstanding >= 1
            rc, loss = self._fc.loss_lagged(post, take, h, self.stream, want)
            if rc != 0:
                _lib.check(rc)
            if not want:
                self._loss_outstanding += 1
            return loss
        self.loss_post()
        return self.loss_take() if self._loss_outstanding > 1 else None

    def hvp(self, direction):
        """H . direction at the batch of the last update (dqnb_learner_hvp); flat float32 in, flat float32 out."""
        v = np.ascontiguousarray(direction, dtype=np.float32).ravel()
        if v.size != self.num_params:
            raise ValueError("direction has %d entries, the learner %d parameters" % (v.size, self.num_params))
        dv, do = ctypes.c_void_p(), ctypes.c_void_p()
        _lib.call("dqnb_dev_alloc", self.device, v.nbytes, ctypes.byref(dv))
        _lib.call("dqnb_dev_alloc", self.device, v.nbytes, ctypes.byref(do))
        try:
            _lib.call("dqnb_copy_h2d", dv, v.ctypes.data, v.nbytes, self.stream)
            _lib.call("dqnb_learner_hvp", self._h, dv, do, self.stream)
            out = np.empty_like(v)
            _lib.call("dqnb_copy_d2h", out.ctypes.data, do, out.nbytes, self.stream)
        finally:
            _lib.call("dqnb_dev_free", self.device, dv)
            _lib.call("dqnb_dev_free", self.device, do)
        return out

    def set_row_weights(self, weights):
        """Host write of DQNB_ROW_WEIGHT (synchronous): the weights the next update(s) will use."""
        w = np.ascontiguousarray(weights, dtype=np.float32)
        _lib.call("dqnb_learner_write", self._h, ROW_WEIGHT, w.ctypes.data, w.size, self.stream)

    def update_host(self, s_tm1, a_tm1, r_t, discount_t, s_t):
        """`update(transitions)` with NumPy inputs, like the reference's jit dispatch."""
        s0 = np.ascontiguousarray(s_tm1, dtype=np.uint8)
        s1 = np.ascontiguousarray(s_t, dtype=np.uint8)
        a = np.ascontiguousarray(a_tm1, dtype=np.int32)
        r = np.ascontiguousarray(r_t, dtype=np.float32)
        d = np.ascontiguousarray(discount_t, dtype=np.float32)
        want = (self.batch_size, 84, 84, 4)
        if s0.shape != want or s1.shape != want or a.shape != (self.batch_size,):
            raise ValueError("batch shapes %s %s %s do not match batch_size=%d"
                             % (s0.shape, s1.shape, a.shape, self.batch_size))
        _lib.call("dqnb_learner_update_host", self._h, s0.ctypes.data, a.ctypes.data, r.ctypes.data,
                  d.ctypes.data, s1.ctypes.data, self.stream)

    def update_device(self, s_tm1_ptr, a_ptr, r_ptr, d_ptr, s_t_ptr):
        _lib.call("dqnb_learner_update", self._h, s_tm1_ptr, a_ptr, r_ptr, d_ptr, s_t_ptr, self.stream)

    def update_from_replay(self, frame_store, slots, weights=None):
        """`weights`: per-row weights of this batch (row_weighted learners; prioritized replay's importance weights)."""
        if weights is None and self._fc is not None:
            fn, h = self._fc_update
            rc = self._fc.update(fn, h, frame_store.handle.value, slots, self.stream, self.batch_size)
            if rc == 0:
                return
            if rc != -1000:
                _lib.check(rc)
        slots = np.ascontiguousarray(slots, dtype=np.int64)
        if slots.size != self.batch_size:
            raise ValueError("need %d slots, got %d" % (self.batch_size, slots.size))
        if weights is None:
            _lib.check(self._update_from_replay(self._h, frame_store.handle, slots.ctypes.data, self.stream))
            return
        w = np.ascontiguousarray(weights, dtype=np.float32)
        if w.shape != (self.batch_size,):
            raise ValueError("need %d weights, got %s" % (self.batch_size, w.shape))
        _lib.call("dqnb_learner_update_from_replay_weighted", self._h, frame_store.handle, slots.ctypes.data,
                  w.ctypes.data, self.stream)

    def sync_target(self):
        _lib.call("dqnb_learner_sync_target", self._h, self.stream)

    def q_values(self, observations, which=ONLINE):
        s = np.ascontiguousarray(observations, dtype=np.uint8)
        if s.ndim == 3:
            s = s[None]
        q = np.empty((s.shape[0], self.num_actions), dtype=np.float32)
        _lib.call("dqnb_learner_q_values", self._h, which, s.ctypes.data, s.shape[0], q.ctypes.data,
                  self.stream)
        return q

    def profile(self, iters=20):
        """Per-kernel device times (us) of one update; advances the state by `iters` updates."""
        buf = ctypes.create_string_buffer(8192)
        _lib.call("dqnb_learner_profile", self._h, int(iters), buf, len(buf), self.stream)
        return buf.value.decode()

    def timeline(self):
        """Begin/end (us) of every kernel of one graph update; advances the state by 3 updates."""
        buf = ctypes.create_string_buffer(8192)
        _lib.call("dqnb_learner_timeline", self._h, buf, len(buf), self.stream)
        return buf.value.decode()

    def launches_per_update(self):
        n = ctypes.c_int64()
        _lib.call("dqnb_learner_launches_per_update", self._h, ctypes.byref(n))
        return n.value

    def synchronize(self):
        _lib.call("dqnb_stream_sync", self.stream)


class MgscConfig(ctypes.Structure):
    _fields_ = [("meta_batch_size", ctypes.c_int32), ("learning_rate", ctypes.c_float),
                ("b1", ctypes.c_float), ("b2", ctypes.c_float), ("eps", ctypes.c_float),
                ("second_order", ctypes.c_int32)]


MGSC_ADAM_M, MGSC_ADAM_V, MGSC_P, MGSC_DL_DP, MGSC_G, MGSC_V, MGSC_THETA_EXP = range(7)


class MetaLearner:
    """`meta_update` of dqn_mgsc_batched/agent.py:160-220 over libdqnb200 (dqnb_mgsc_*).

    Reads the learner's params / RMSProp state (virtual steps only), owns the optax.adam state of
    the (meta_batch_size,) logits vector.
    """

    def __init__(self, learner: Learner, meta_batch_size, learning_rate=0.00025, b1=0.9, b2=0.999, eps=1e-8,
                 second_order=False):
        if learner.double_q:
            raise ValueError("MGSC uses the q_learning loss: the learner must have double_q=False")
        self.learner, self.meta_batch_size = learner, int(meta_batch_size)
        self.second_order = bool(second_order)
        cfg = MgscConfig(self.meta_batch_size, learning_rate, b1, b2, eps, int(self.second_order))
        h = ctypes.c_void_p()
        _lib.call("dqnb_mgsc_create", learner._h, ctypes.byref(cfg), ctypes.byref(h))
        self._h = h

    def close(self):
        if getattr(self, "_h", None) is not None:
            _lib.call("dqnb_mgsc_destroy", self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @staticmethod
    def _online(transition):
        s0 = np.ascontiguousarray(transition[0], dtype=np.uint8)
        s1 = np.ascontiguousarray(transition[4], dtype=np.uint8)
        if s0.shape != (84, 84, 4) or s1.shape != (84, 84, 4):
            raise ValueError("online transition stacks must be (84, 84, 4) uint8, got %s / %s" % (s0.shape, s1.shape))
        return s0, int(transition[1]), float(transition[2]), float(transition[3]), s1

    def meta_update(self, transitions, logits, online_transition):
        """Host meta batch (stacked NumPy fields).  Returns (new_logits float32 [M], meta_loss)."""
        M = self.meta_batch_size
        s0 = np.ascontiguousarray(transitions[0], dtype=np.uint8)
        s1 = np.ascontiguousarray(transitions[4], dtype=np.uint8)
        a = np.ascontiguousarray(transitions[1], dtype=np.int32)
        r = np.ascontiguousarray(transitions[2], dtype=np.float32)
        d = np.ascontiguousarray(transitions[3], dtype=np.float32)
        lg = np.ascontiguousarray(logits, dtype=np.float32)
        if s0.shape != (M, 84, 84, 4) or s1.shape != s0.shape or a.shape != (M,) or lg.shape != (M,):
            raise ValueError("meta batch shapes %s %s %s do not match meta_batch_size=%d"
                             % (s0.shape, a.shape, lg.shape, M))
        o0, oa, orr, od, o1 = self._online(online_transition)
        out, loss = np.empty(M, dtype=np.float32), ctypes.c_float()
        _lib.call("dqnb_mgsc_meta_update_host", self._h, s0.ctypes.data, a.ctypes.data, r.ctypes.data,
                  d.ctypes.data, s1.ctypes.data, lg.ctypes.data, o0.ctypes.data, oa, orr, od, o1.ctypes.data,
                  out.ctypes.data, ctypes.byref(loss), self.learner.stream)
        return out, float(loss.value)

    def meta_update_from_replay(self, frame_store, slots, logits, online_transition):
        """Meta batch gathered on the device from replay slots."""
        M = self.meta_batch_size
        sl = np.ascontiguousarray(slots, dtype=np.int64)
        lg = np.ascontiguousarray(logits, dtype=np.float32)
        if sl.shape != (M,) or lg.shape != (M,):
            raise ValueError("need %d slots and logits, got %s / %s" % (M, sl.shape, lg.shape))
        o0, oa, orr, od, o1 = self._online(online_transition)
        out, loss = np.empty(M, dtype=np.float32), ctypes.c_float()
        _lib.call("dqnb_mgsc_meta_update", self._h, frame_store.handle, sl.ctypes.data, lg.ctypes.data,
                  o0.ctypes.data, oa, orr, od, o1.ctypes.data, out.ctypes.data, ctypes.byref(loss),
                  self.learner.stream)
        return out, float(loss.value)

    def read(self, which):
        n = self.meta_batch_size if which <= MGSC_DL_DP else self.learner.num_params
        out = np.empty(n, dtype=np.float32)
        _lib.call("dqnb_mgsc_read", self._h, which, out.ctypes.data, n)
        return out

    def read_learner(self, which_learner, which, n):
        """Test hook: array `which` (n floats) of the internal batch-of-M (0) / batch-of-1 (1) learner."""
        out = np.empty(n, dtype=np.float32)
        _lib.call("dqnb_mgsc_read_learner", self._h, which_learner, which, out.ctypes.data, n)
        return out

    def get_state(self):
        """(count, mu, nu) like optax ScaleByAdamState."""
        c = ctypes.c_int32()
        _lib.call("dqnb_mgsc_adam_count", self._h, ctypes.byref(c), None)
        return {"count": int(c.value), "mu": self.read(MGSC_ADAM_M), "nu": self.read(MGSC_ADAM_V)}

    def set_state(self, state):
        for which, key in ((MGSC_ADAM_M, "mu"), (MGSC_ADAM_V, "nu")):
            v = np.ascontiguousarray(state[key], dtype=np.float32)
            _lib.call("dqnb_mgsc_write", self._h, which, v.ctypes.data, v.size)
        c = ctypes.c_int32(int(state["count"]))
        _lib.call("dqnb_mgsc_adam_count", self._h, None, ctypes.byref(c))
