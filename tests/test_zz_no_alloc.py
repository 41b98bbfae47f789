"""dqnb_logsumexp_noalloc (the allocation-free entry point: scratch in a file-scope __device__ array, address
taken with cudaGetSymbolAddress) against the existing dqnb_logits_lse on the same logits, through the C ABI.

Kept in a file that pytest collects LAST: a failure here must not hide the
other tests.  Passes alone in a fresh process and after the rest of the suite (it shares no state with it: the
static scratch array is only ever touched by this entry point)."""
import ctypes

import numpy as np
import pytest

from dqn_mgsc_zoo_b200 import _lib

pytestmark = pytest.mark.gpu
ATOL = 2e-6          # the logit tests' tolerance (tests/test_mgsc_gpu.py); same kernels => equal bits in practice
MAX_LOGITS = 4_194_304


class DeviceBuffer:
    """A raw device buffer from the C ABI's own allocator (the TEST may allocate; the entry point may not)."""

    def __init__(self, nbytes, device=0):
        self.device, self.ptr = device, ctypes.c_void_p()
        _lib.call("dqnb_dev_alloc", device, int(nbytes), ctypes.byref(self.ptr))

    def upload(self, array):
        a = np.ascontiguousarray(array)
        _lib.call("dqnb_copy_h2d", self.ptr, a.ctypes.data, a.nbytes, 0)
        _lib.call("dqnb_stream_sync", 0)

    def download(self, dtype, count):
        out = np.empty(count, dtype=dtype)
        _lib.call("dqnb_copy_d2h", out.ctypes.data, self.ptr, out.nbytes, 0)
        return out

    def free(self):
        _lib.call("dqnb_dev_free", self.device, self.ptr)


def lse_existing(logits):
    """The existing entry point: a logit ring handle holding `logits`, dqnb_logits_lse."""
    h = ctypes.c_void_p()
    _lib.call("dqnb_logits_create", logits.size, 0, ctypes.byref(h))
    try:
        _lib.call("dqnb_logits_write", h, logits.ctypes.data, logits.size)
        out = ctypes.c_float()
        _lib.call("dqnb_logits_lse", h, ctypes.byref(out), 0)
        return np.float32(out.value)
    finally:
        _lib.call("dqnb_logits_destroy", h)


def test_noalloc_logsumexp_matches_logits_lse_on_two_calls_in_a_row():
    rng = np.random.default_rng(5)
    # the first call is LONGER than the second: stale partials of call 1 stay in the shared scratch and must not
    # leak into call 2; sizes off the 1024-logit block boundary; -inf entries like an MGSC ring's empty slots
    first = rng.normal(0.0, 2.0, 300_007).astype(np.float32)
    first[::97] = -np.inf
    second = rng.normal(-3.0, 0.5, 5_003).astype(np.float32)
    second[17] = 11.0                                   # a dominant entry
    src = DeviceBuffer(first.nbytes)
    out = DeviceBuffer(16)
    try:
        got = []
        for logits in (first, second):
            src.upload(logits)
            out.upload(np.full(4, np.nan, np.float32))
            _lib.call("dqnb_logsumexp_noalloc", 0, src.ptr, logits.size, out.ptr, 0)
            _lib.call("dqnb_stream_sync", 0)
            got.append(out.download(np.float32, 1)[0])
        for logits, g in zip((first, second), got):
            want = lse_existing(logits)
            ref = logits.astype(np.float64)
            exact = ref.max() + np.log(np.exp(ref - ref.max()).sum())
            assert np.isfinite(g)
            np.testing.assert_allclose(g, want, rtol=0, atol=ATOL)
            np.testing.assert_allclose(g, exact, rtol=0, atol=2e-5)     # and both are right, not merely equal
        # a third call repeating the first input gives the first answer again (the shared scratch carries nothing over)
        src.upload(first)
        _lib.call("dqnb_logsumexp_noalloc", 0, src.ptr, first.size, out.ptr, 0)
        _lib.call("dqnb_stream_sync", 0)
        assert out.download(np.float32, 1)[0] == got[0]
    finally:
        src.free()
        out.free()


def test_noalloc_logsumexp_rejects_sizes_beyond_its_fixed_scratch():
    buf = DeviceBuffer(64)
    try:
        with pytest.raises(_lib.DqnbError, match="allocation-free limit"):
            _lib.call("dqnb_logsumexp_noalloc", 0, buf.ptr, MAX_LOGITS + 1, buf.ptr, 0)
        with pytest.raises(_lib.DqnbError, match="at least one"):
            _lib.call("dqnb_logsumexp_noalloc", 0, buf.ptr, 0, buf.ptr, 0)
        with pytest.raises(_lib.DqnbError, match="null"):
            _lib.call("dqnb_logsumexp_noalloc", 0, None, 8, buf.ptr, 0)
    finally:
        buf.free()
