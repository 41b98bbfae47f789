"""ctypes binding of libdqnb200.so (the C ABI declared in include/dqnb200.h).

There is no CPU fallback: if the shared library is missing or a call fails, an
exception is raised.  Build the library with `python -m dqn_mgsc_zoo_b200.build`
(or `__graft_entry__.build()`).
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libdqnb200.so")

c_i64, c_i32, c_int, c_f32 = ctypes.c_int64, ctypes.c_int32, ctypes.c_int, ctypes.c_float
c_vp, c_up, c_sz = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_size_t
STREAM = ctypes.c_size_t   # uintptr_t

# name -> argtypes; every function returns int status except dqnb_version/dqnb_last_error.
SIGNATURES = {
    "dqnb_device_count": [ctypes.POINTER(c_int)],
    "dqnb_dev_alloc": [c_int, c_sz, ctypes.POINTER(c_vp)],
    "dqnb_dev_free": [c_int, c_vp],
    "dqnb_copy_h2d": [c_vp, c_vp, c_sz, STREAM],
    "dqnb_copy_d2h": [c_vp, c_vp, c_sz, STREAM],
    "dqnb_stream_sync": [STREAM],
    "dqnb_replay_create": [c_i64, c_int, c_int, c_int, c_i64, c_int, ctypes.POINTER(c_vp)],
    "dqnb_replay_destroy": [c_vp],
    "dqnb_replay_put": [c_vp, c_i64, c_vp, c_vp, c_i32, c_f32, c_f32, STREAM],
    "dqnb_replay_flush": [c_vp, STREAM],
    "dqnb_replay_gather": [c_vp, c_vp, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, STREAM],
    "dqnb_replay_gather_dev": [c_vp, c_vp, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, STREAM],
    "dqnb_replay_gather_host": [c_vp, c_vp, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, STREAM],
    "dqnb_replay_set_gather_splits": [c_vp, c_int],
    "dqnb_replay_stats": [c_vp, ctypes.POINTER(c_i64)],
    "dqnb_replay_export_size": [c_vp, ctypes.POINTER(c_i64)],
    "dqnb_replay_export": [c_vp, c_vp, c_i64, STREAM],
    "dqnb_replay_import": [c_vp, c_vp, c_i64, STREAM],
    "dqnb_learner_create": [c_vp, c_int, ctypes.POINTER(c_vp)],
    "dqnb_learner_destroy": [c_vp],
    "dqnb_learner_num_params": [c_vp, ctypes.POINTER(c_i64)],
    "dqnb_learner_write": [c_vp, c_int, c_vp, c_i64, STREAM],
    "dqnb_learner_read": [c_vp, c_int, c_vp, c_i64, STREAM],
    "dqnb_learner_update": [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, STREAM],
    "dqnb_learner_update_host": [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, STREAM],
    "dqnb_learner_update_from_replay": [c_vp, c_vp, c_vp, STREAM],
    "dqnb_learner_update_from_replay_weighted": [c_vp, c_vp, c_vp, c_vp, STREAM],
    "dqnb_learner_sync_target": [c_vp, STREAM],
    "dqnb_learner_q_values": [c_vp, c_int, c_vp, c_i32, c_vp, STREAM],
    "dqnb_learner_device_ptr": [c_vp, c_int, ctypes.POINTER(c_vp), ctypes.POINTER(c_i64)],
    "dqnb_learner_write_dev": [c_vp, c_int, c_vp, c_i64, STREAM],
    "dqnb_learner_apply_grads": [c_vp, STREAM],
    "dqnb_learner_adjoint_dots": [c_vp, c_vp, c_vp, STREAM],
    "dqnb_learner_hvp": [c_vp, c_vp, c_vp, STREAM],
    "dqnb_dp_unique_id": [c_vp],
    "dqnb_dp_init": [c_vp, c_vp, c_int, c_int],
    "dqnb_dp_allreduce": [c_vp, STREAM],
    "dqnb_dp_info": [c_vp, ctypes.POINTER(c_int), ctypes.POINTER(c_int), ctypes.POINTER(c_int)],
    "dqnb_learner_get_config": [c_vp, c_vp, ctypes.POINTER(c_int)],
    "dqnb_mgsc_create": [c_vp, c_vp, ctypes.POINTER(c_vp)],
    "dqnb_mgsc_destroy": [c_vp],
    "dqnb_mgsc_meta_update": [c_vp, c_vp, c_vp, c_vp, c_vp, c_i32, c_f32, c_f32, c_vp, c_vp,
                              ctypes.POINTER(c_f32), STREAM],
    "dqnb_mgsc_meta_update_host": [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_i32, c_f32, c_f32, c_vp,
                                   c_vp, ctypes.POINTER(c_f32), STREAM],
    "dqnb_mgsc_read": [c_vp, c_int, c_vp, c_i64],
    "dqnb_mgsc_write": [c_vp, c_int, c_vp, c_i64],
    "dqnb_mgsc_read_learner": [c_vp, c_int, c_int, c_vp, c_i64],
    "dqnb_mgsc_adam_count": [c_vp, ctypes.POINTER(c_i32), ctypes.POINTER(c_i32)],
    "dqnb_logits_create": [c_i64, c_int, ctypes.POINTER(c_vp)],
    "dqnb_logits_destroy": [c_vp],
    "dqnb_logits_fill": [c_vp, c_i64, c_i64, STREAM],
    "dqnb_logits_set": [c_vp, c_vp, c_vp, c_i64, STREAM],
    "dqnb_logits_get": [c_vp, c_vp, c_vp, c_i64, STREAM],
    "dqnb_logits_sample": [c_vp, c_vp, c_i64, c_vp, STREAM],
    "dqnb_logits_lse": [c_vp, ctypes.POINTER(c_f32), STREAM],
    "dqnb_logits_read": [c_vp, c_vp, c_i64],
    "dqnb_logits_write": [c_vp, c_vp, c_i64],
    "dqnb_logsumexp_noalloc": [c_int, c_vp, c_i64, c_vp, STREAM],
    "dqnb_sumtree_create": [c_i64, c_int, ctypes.POINTER(c_vp)],
    "dqnb_sumtree_destroy": [c_vp],
    "dqnb_sumtree_set": [c_vp, c_vp, c_vp, c_i64, STREAM],
    "dqnb_sumtree_query": [c_vp, c_vp, c_i64, c_int, c_vp, ctypes.POINTER(ctypes.c_double), STREAM],
    "dqnb_sumtree_read": [c_vp, c_vp, c_i64],
    "dqnb_sumtree_write": [c_vp, c_vp, c_i64],
    "dqnb_learner_loss_post": [c_vp, STREAM],
    "dqnb_learner_loss_take": [c_vp, ctypes.POINTER(ctypes.c_float)],
    "dqnb_learner_launches_per_update": [c_vp, ctypes.POINTER(c_i64)],
    "dqnb_learner_profile": [c_vp, c_int, ctypes.c_char_p, c_sz, STREAM],
    "dqnb_learner_timeline": [c_vp, ctypes.c_char_p, c_sz, STREAM],
    "dqnb_learner_relaunch": [c_vp, c_int, STREAM],
    "dqnb_selftest_tf32_gemm": [c_vp, c_vp, c_vp, c_int, c_int, c_int, c_int, c_int, c_int, c_int, STREAM],
}


class DqnbError(RuntimeError):
    pass


_lib = None


def load():
    """Loads libdqnb200.so (once) and declares every signature.  Raises if absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise DqnbError(
            "libdqnb200.so not found at %s - build it with `python -m dqn_mgsc_zoo_b200.build`; "
            "there is no CPU fallback for the learner hot path" % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    lib.dqnb_version.restype = c_int
    lib.dqnb_version.argtypes = []
    lib.dqnb_last_error.restype = ctypes.c_char_p
    lib.dqnb_last_error.argtypes = []
    for name, argtypes in SIGNATURES.items():
        fn = getattr(lib, name)       # AttributeError if the symbol is missing
        fn.restype = c_int
        fn.argtypes = argtypes
    _lib = lib
    return lib


def check(status):
    if status != 0:
        raise DqnbError("libdqnb200 error %d: %s" % (status, load().dqnb_last_error().decode()))


def call(name, *args):
    check(getattr(load(), name)(*args))


def fastcall():
    """The optional CPython helper for the per-step calls (csrc_host/fastcall.c), or None when it is not built.
    It calls the same C-ABI entry points as the ctypes wrappers; only the per-call Python overhead differs."""
    global _FASTCALL
    try:
        return _FASTCALL
    except NameError:
        pass
    try:
        from dqn_mgsc_zoo_b200 import _fastcall as mod
    except ImportError:
        mod = None
    if os.environ.get("DQNB_NO_FASTCALL"):
        mod = None
    _FASTCALL = mod
    return mod


def fn_address(name):
    """Address of an exported C-ABI function, for the helper."""
    return ctypes.cast(getattr(load(), name), ctypes.c_void_p).value
