"""The optional CPython helper for the per-step calls (dqn_mgsc_zoo_b200/csrc_host/fastcall.c) against a stub
library with the C-ABI signatures of dqnb_replay_put / dqnb_learner_update_from_replay / loss_post / loss_take:
what arrives at the C entry point must be exactly what the ctypes path would pass, and anything the helper does
not handle must come back as FALLBACK (the caller then takes the ctypes path).  No GPU."""
import ctypes
import os
import subprocess
import sysconfig

import numpy as np
import pytest

from dqn_mgsc_zoo_b200 import build as builder

STUB = r"""
#include <stdint.h>
#include <string.h>
static struct { void* h; int64_t slot; uint8_t first0, last0, first1, last1; int32_t action; float reward, discount;
                uintptr_t stream; int64_t slots[4]; void* rep; int posts, takes; } rec;
int stub_put(void* h, int64_t slot, const uint8_t* a, const uint8_t* b, int32_t action, float reward, float discount,
             uintptr_t stream) {
  rec.h = h; rec.slot = slot; rec.first0 = a[0]; rec.last0 = a[84 * 84 * 4 - 1]; rec.first1 = b[0];
  rec.last1 = b[84 * 84 * 4 - 1]; rec.action = action; rec.reward = reward; rec.discount = discount; rec.stream = stream;
  return slot == 77 ? 5 : 0;
}
int stub_update(void* h, void* rep, const int64_t* slots, uintptr_t stream) {
  rec.h = h; rec.rep = rep; memcpy(rec.slots, slots, sizeof(rec.slots)); rec.stream = stream; return 0;
}
int stub_post(void* h, uintptr_t stream) { (void)h; (void)stream; rec.posts++; return 0; }
int stub_take(void* h, float* loss) { (void)h; rec.takes++; *loss = 1.5f; return 0; }
void* stub_rec(void) { return &rec; }
"""


class Rec(ctypes.Structure):
    _fields_ = [("h", ctypes.c_void_p), ("slot", ctypes.c_int64), ("first0", ctypes.c_uint8), ("last0", ctypes.c_uint8),
                ("first1", ctypes.c_uint8), ("last1", ctypes.c_uint8), ("action", ctypes.c_int32),
                ("reward", ctypes.c_float), ("discount", ctypes.c_float), ("stream", ctypes.c_size_t),
                ("slots", ctypes.c_int64 * 4), ("rep", ctypes.c_void_p), ("posts", ctypes.c_int), ("takes", ctypes.c_int)]


@pytest.fixture(scope="module")
def env(tmp_path_factory):
    builder.build_fastcall()
    from dqn_mgsc_zoo_b200 import _fastcall
    d = tmp_path_factory.mktemp("stub")
    src, so = os.path.join(d, "stub.c"), os.path.join(d, "stub.so")
    with open(src, "w") as f:
        f.write(STUB)
    subprocess.run([os.environ.get("CC", "gcc"), "-O1", "-shared", "-fPIC", "-o", so, src], check=True)
    lib = ctypes.CDLL(so)
    lib.stub_rec.restype = ctypes.POINTER(Rec)
    addr = lambda name: ctypes.cast(getattr(lib, name), ctypes.c_void_p).value
    return _fastcall, lib, addr


def test_put_passes_buffers_and_numpy_scalars(env):
    fc, lib, addr = env
    s0 = np.arange(84 * 84 * 4, dtype=np.uint8).reshape(84, 84, 4)
    s1 = (s0 + 3).astype(np.uint8)
    s1.flags.writeable = False                      # read-only observations are fine
    rc = fc.put(addr("stub_put"), 0x1234, np.int64(9), s0, s1, np.int32(4), np.float32(0.25), np.float64(0.99), 7,
                84, 84, 4)
    r = lib.stub_rec().contents
    assert rc == 0 and r.h == 0x1234 and r.slot == 9 and r.action == 4 and r.stream == 7
    assert (r.first0, r.last0, r.first1, r.last1) == (s0.flat[0], s0.flat[-1], s1.flat[0], s1.flat[-1])
    assert r.reward == np.float32(0.25) and r.discount == np.float32(0.99)
    assert fc.put(addr("stub_put"), 1, 77, s0, s0, 0, 0.0, 1.0, 0, 84, 84, 4) == 5    # C-ABI status passes through


def test_put_falls_back_on_what_it_does_not_handle(env):
    fc, _, addr = env
    ok = np.zeros((84, 84, 4), np.uint8)
    for bad in (np.zeros((84, 84, 4), np.float32), np.zeros((84, 4, 84), np.uint8), np.zeros((84, 84, 8), np.uint8)[:, :, ::2],
                [[0]], None):
        assert fc.put(addr("stub_put"), 1, 0, bad, ok, 0, 0.0, 1.0, 0, 84, 84, 4) == fc.FALLBACK
        assert fc.put(addr("stub_put"), 1, 0, ok, bad, 0, 0.0, 1.0, 0, 84, 84, 4) == fc.FALLBACK
    assert fc.put(addr("stub_put"), 1, 0, ok, ok, "x", 0.0, 1.0, 0, 84, 84, 4) == fc.FALLBACK


def test_update_and_loss_lagged(env):
    fc, lib, addr = env
    slots = np.array([5, 6, 7, 1 << 40], dtype=np.int64)
    assert fc.update(addr("stub_update"), 11, 22, slots, 3, 4) == 0
    r = lib.stub_rec().contents
    assert (r.h, r.rep, r.stream, list(r.slots)) == (11, 22, 3, slots.tolist())
    assert fc.update(addr("stub_update"), 11, 22, slots.astype(np.int32), 3, 4) == fc.FALLBACK   # dtype
    assert fc.update(addr("stub_update"), 11, 22, slots[:3], 3, 4) == fc.FALLBACK                 # length
    assert fc.update(addr("stub_update"), 11, 22, slots.tolist(), 3, 4) == fc.FALLBACK            # not a buffer
    p0, t0 = r.posts, r.takes
    assert fc.loss_lagged(addr("stub_post"), addr("stub_take"), 11, 0, False) == (0, None)
    assert fc.loss_lagged(addr("stub_post"), addr("stub_take"), 11, 0, True) == (0, 1.5)
    r = lib.stub_rec().contents
    assert (r.posts - p0, r.takes - t0) == (2, 1)
