"""TEST INFRASTRUCTURE ONLY — loads the *unmodified* reference replay modules.

`/root/reference/dqn_zoo/replay.py` and `replay_circular.py` import jax, snappy,
dm_env and `dqn_zoo.parts` at module top (replay.py:24-31) but their
FIFO / reservoir / circular / MGSC classes only need numpy + collections at run
time.  This shim injects stub modules and loads the two files straight from the
read-only reference tree.  It exists only in the build container (the GPU box
has no /root/reference): it is used by `tests/golden/make_golden.py` to
generate committed fixtures and by CPU tests (skipped when the tree is absent)
to pin `oracle/replay_oracle.py` against the real reference.

Nothing in the product package may import this file.
"""
import collections
import importlib.util
import os
import sys
import types

import numpy as np

# The read-only reference tree in the build container, else the byte-identical copies oracle/build_ref.py
# staged under the git-ignored oracle/_ref/ (they travel to the GPU box with the snapshot).
_CANDIDATES = (os.environ.get("DQNB_REFERENCE_ROOT", "/root/reference"),
               os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref"))


def _root():
    for root in _CANDIDATES:
        if os.path.isfile(os.path.join(root, "dqn_zoo", "replay.py")):
            return root
    return None


REFERENCE_ROOT = _root() or _CANDIDATES[0]


def available() -> bool:
    return _root() is not None


def _stub_modules():
    jnp = types.ModuleType("jax.numpy")
    jnp.ndarray = np.ndarray
    jnp.DeviceArray = type("DeviceArray", (), {})
    for name in ("isnan", "isfinite", "exp", "log", "sum", "asarray", "zeros"):
        setattr(jnp, name, getattr(np, name))
    jax = types.ModuleType("jax")
    jax.numpy = jnp
    jax.tree_util = types.SimpleNamespace(
        tree_flatten=lambda t: (list(t) if isinstance(t, (list, tuple)) else [t], None))
    snappy = types.ModuleType("snappy")
    snappy.compress = lambda a: np.ascontiguousarray(a).tobytes()
    snappy.uncompress = lambda b: b
    dm_env = types.ModuleType("dm_env")
    TimeStep = collections.namedtuple("TimeStep", "step_type reward discount observation")
    TimeStep.first = lambda self: self.step_type == 0
    TimeStep.mid = lambda self: self.step_type == 1
    TimeStep.last = lambda self: self.step_type == 2
    dm_env.TimeStep = TimeStep
    dm_env.StepType = types.SimpleNamespace(FIRST=0, MID=1, LAST=2)
    dm_env.Environment = object
    pkg = types.ModuleType("dqn_zoo")
    pkg.__path__ = []
    parts = types.ModuleType("dqn_zoo.parts")
    parts.Action = int
    pkg.parts = parts
    return {"jax": jax, "jax.numpy": jnp, "snappy": snappy, "dm_env": dm_env,
            "dqn_zoo": pkg, "dqn_zoo.parts": parts}


_cache = {}


def load(name: str):
    """Returns the reference module `dqn_zoo/<name>.py` ('replay', 'replay_circular')."""
    if name in _cache:
        return _cache[name]
    if not available():
        raise FileNotFoundError("reference tree not present at %s" % (_CANDIDATES,))
    stubs = _stub_modules()
    saved = {k: sys.modules.get(k) for k in stubs}
    sys.modules.update(stubs)
    try:
        path = os.path.join(_root(), "dqn_zoo", name + ".py")
        spec = importlib.util.spec_from_file_location("_dqnref_" + name, path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    mod.TimeStep = stubs["dm_env"].TimeStep
    _cache[name] = mod
    return mod
