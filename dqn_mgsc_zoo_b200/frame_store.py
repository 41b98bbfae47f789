"""Thin Python handle over the C-ABI replay frame store (dqnb_replay_*)."""
import ctypes

import numpy as np

from dqn_mgsc_zoo_b200 import _lib


def _ptr(arr):
    return arr.ctypes.data


_addressof, _from_buffer = ctypes.addressof, ctypes.c_char.from_buffer


def _fast_ptr(arr):
    """Address of a C-contiguous array's data through the buffer protocol: ~1 us instead of the ~2 us of
    `arr.ctypes.data` (which builds a ctypes helper object); read-only arrays take the slow road."""
    try:
        return _addressof(_from_buffer(arr))
    except TypeError:
        return arr.ctypes.data


class FrameStore:
    """uint8 frame planes + per-slot (action, reward, discount) resident in HBM."""

    def __init__(self, capacity, height, width, stack, device=0, stream=0, initial_planes=0):
        self.capacity, self.shape, self.device, self.stream = int(capacity), (height, width, stack), device, stream
        h = ctypes.c_void_p()
        _lib.call("dqnb_replay_create", self.capacity, height, width, stack, int(initial_planes), device,
                  ctypes.byref(h))
        self._h = h
        self._put = _lib.load().dqnb_replay_put
        self._last = (None, 0)
        fc = _lib.fastcall()
        self._fc_put = fc.put if fc is not None else None
        self._fc_args = (_lib.fn_address("dqnb_replay_put"), h.value)

    @property
    def handle(self):
        return self._h

    def close(self):
        if self._h is not None:
            _lib.call("dqnb_replay_destroy", self._h)
            self._h = None
            self._fc_put = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def put(self, slot, s_tm1, s_t, action, reward, discount):
        fc = self._fc_put
        if fc is not None:   # buffer-protocol fast path: same C entry point, ~3 us less Python per call
            fn, h = self._fc_args
            hh, ww, ss = self.shape
            rc = fc(fn, h, slot, s_tm1, s_t, action, reward, discount, self.stream, hh, ww, ss)
            if rc == 0:
                return
            if rc != -1000:
                _lib.check(rc)
        s_tm1 = np.ascontiguousarray(s_tm1, dtype=np.uint8)
        s_t = np.ascontiguousarray(s_t, dtype=np.uint8)
        if s_tm1.shape != self.shape or s_t.shape != self.shape:
            raise ValueError("stack shape %s/%s does not match store %s" % (s_tm1.shape, s_t.shape, self.shape))
        # `arr.ctypes.data` costs ~1-2 us; the accumulator hands over the previous s_t object as the
        # next s_tm1 (replay.py:1184-1190), so its address is remembered (the array is kept alive here).
        last = self._last
        p0 = last[1] if last[0] is s_tm1 else _fast_ptr(s_tm1)
        p1 = _fast_ptr(s_t)
        self._last = (s_t, p1)
        _lib.check(self._put(self._h, int(slot), p0, p1, int(action), float(reward), float(discount), self.stream))

    def flush(self):
        _lib.call("dqnb_replay_flush", self._h, self.stream)

    def gather_host(self, slots):
        """Returns (s_tm1, a_i32, r_f32, disc_f32, s_t) NumPy arrays for host int64 slots."""
        slots = np.ascontiguousarray(slots, dtype=np.int64)
        n = slots.size
        s1 = np.empty((n,) + self.shape, dtype=np.uint8)
        s2 = np.empty((n,) + self.shape, dtype=np.uint8)
        a = np.empty(n, dtype=np.int32)
        r = np.empty(n, dtype=np.float32)
        d = np.empty(n, dtype=np.float32)
        _lib.call("dqnb_replay_gather_host", self._h, _ptr(slots), n, _ptr(s1), _ptr(s2), _ptr(a), _ptr(r),
                  _ptr(d), self.stream)
        return s1, a, r, d, s2

    def gather(self, slots, s_tm1_ptr, s_t_ptr, a_ptr, r_ptr, d_ptr, stream=None):
        """Gather into caller-provided device pointers (ints), asynchronously."""
        slots = np.ascontiguousarray(slots, dtype=np.int64)
        _lib.call("dqnb_replay_gather", self._h, _ptr(slots), slots.size, s_tm1_ptr, s_t_ptr, a_ptr, r_ptr,
                  d_ptr, self.stream if stream is None else stream)

    def gather_dev(self, slots_dev_ptr, n, s_tm1_ptr, s_t_ptr, a_ptr, r_ptr, d_ptr, stream=None):
        """Same with the int32 slot list already in device memory (no upload): the bare kernel."""
        _lib.call("dqnb_replay_gather_dev", self._h, slots_dev_ptr, int(n), s_tm1_ptr, s_t_ptr, a_ptr, r_ptr,
                  d_ptr, self.stream if stream is None else stream)

    def export_blob(self):
        """The whole store as one uint8 array (dqnb_replay_export: tables + plane pool, one D2H pass)."""
        n = ctypes.c_int64()
        _lib.call("dqnb_replay_export_size", self._h, ctypes.byref(n))
        blob = np.empty(n.value, dtype=np.uint8)
        # the size depends on the free list, which a flush inside export may grow: retry once with the new size
        try:
            _lib.call("dqnb_replay_export", self._h, blob.ctypes.data, blob.size, self.stream)
        except _lib.DqnbError:
            _lib.call("dqnb_replay_export_size", self._h, ctypes.byref(n))
            blob = np.empty(n.value, dtype=np.uint8)
            _lib.call("dqnb_replay_export", self._h, blob.ctypes.data, blob.size, self.stream)
        return blob

    def import_blob(self, blob):
        blob = np.ascontiguousarray(blob, dtype=np.uint8)
        _lib.call("dqnb_replay_import", self._h, blob.ctypes.data, blob.size, self.stream)
        self._last = (None, 0)

    def set_gather_splits(self, splits):
        _lib.call("dqnb_replay_set_gather_splits", self._h, int(splits))

    def stats(self):
        out = (ctypes.c_int64 * 8)()
        _lib.call("dqnb_replay_stats", self._h, out)
        keys = ("planes_in_use", "pool_planes", "puts", "planes_uploaded", "chained_puts", "pool_growths",
                "hbm_bytes")
        return dict(zip(keys, list(out)))
