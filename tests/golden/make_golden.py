"""Generates tests/golden/replay_golden.npz by running the UNMODIFIED reference
replay classes (loaded from /root/reference through oracle/ref_shim.py) through
the scenarios in tests/replay_scenarios.py.  Run in the build container only:

    python tests/golden/make_golden.py

The reference tree does not travel to the GPU box; the .npz does.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import ref_shim  # noqa: E402
from tests import replay_scenarios as sc  # noqa: E402


def reference_factory(kind):
    r, rc = ref_shim.load("replay"), ref_shim.load("replay_circular")
    cls = {"per": None, "gen_per": None, "fifo": r.TransitionReplay, "reservoir": r.ReservoirTransitionReplay,
           "gen_fifo": rc.TransitionReplay, "gen_reservoir": rc.ReservoirTransitionReplay,
           "mgsc_fifo": rc.MGSCFiFoTransitionReplay,
           "mgsc_reservoir": rc.MGSCReservoirTransitionReplay}[kind]
    if kind in ("per", "gen_per"):
        cls = (r if kind == "per" else rc).PrioritizedTransitionReplay
        return lambda capacity, structure, alpha, beta, usp, normalize, rng: cls(
            capacity, structure, alpha, beta, usp, normalize, rng)
    return lambda capacity, structure, rng: cls(capacity, structure, rng)


def run_reference(name):
    spec = sc.SCENARIOS[name]
    make = reference_factory(spec["kind"])
    fn = getattr(sc, spec["fn"])
    if spec["fn"] == "frames_uniform":
        r = ref_shim.load("replay")
        return fn(make, r.TransitionAccumulator(), r.Transition(None, None, None, None, None),
                  **spec["args"])
    return fn(make, **spec["args"])


def main():
    out = {}
    for name in sc.SCENARIOS:
        for key, val in run_reference(name).items():
            out[name + "/" + key] = np.asarray(val)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "replay_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes,", len(out), "arrays")


if __name__ == "__main__":
    main()
