"""TEST INFRASTRUCTURE ONLY — stages the UNMODIFIED reference replay modules for the GPU box.

`/root/reference` exists only in the build container.  The reference is pure Python (no sources to
compile), so "building" oracle/_ref means copying the two files of the path that run without JAX
(dqn_zoo/replay.py, dqn_zoo/replay_circular.py; see oracle/ref_shim.py for the stubbed imports) byte for
byte into the git-ignored `oracle/_ref/dqn_zoo/`.  Like a built `.so`, that directory is kept out of git
history but travels with the gpurun snapshot, so that on the GPU box

  * `bench.py --impl reference` and the `cpu_baseline` leg time the reference's OWN
    TransitionReplay.add / sample (replay.py:163-243), not a restatement, and
  * the parity tests that compare against the live reference classes also run there.

Called from `__graft_entry__.build()`; a no-op (keeping whatever is staged) when /root/reference is absent.
"""
import hashlib
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
SRC_ROOT = os.environ.get("DQNB_REFERENCE_ROOT", "/root/reference")
DST = os.path.join(HERE, "_ref", "dqn_zoo")
FILES = ("replay.py", "replay_circular.py")


def _sha(path):
    with open(path, "rb") as f:
        return hashlib.sha256(f.read()).hexdigest()


def build(verbose=False):
    """Returns {file: sha256} of what is staged (empty dict if nothing could be staged)."""
    src_dir = os.path.join(SRC_ROOT, "dqn_zoo")
    staged = {}
    if os.path.isdir(src_dir):
        os.makedirs(DST, exist_ok=True)
        for name in FILES:
            src, dst = os.path.join(src_dir, name), os.path.join(DST, name)
            if not os.path.isfile(dst) or _sha(src) != _sha(dst):
                shutil.copyfile(src, dst)
            staged[name] = _sha(dst)
        with open(os.path.join(HERE, "_ref", "MANIFEST"), "w") as f:
            f.write("unmodified copies from %s (sha256):\n" % src_dir)
            for name, digest in staged.items():
                f.write("%s  %s\n" % (digest, name))
    elif os.path.isdir(DST):
        staged = {name: _sha(os.path.join(DST, name)) for name in FILES if os.path.isfile(os.path.join(DST, name))}
    if verbose:
        print("oracle/_ref:", staged or "nothing staged (no reference tree on this machine)")
    return staged


if __name__ == "__main__":
    build(verbose=True)
