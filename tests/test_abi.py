"""CPU checks of the drop-in boundary: the C-ABI library builds, loads and exports every
symbol include/dqnb200.h declares, and the ctypes table covers them all (no compute calls)."""
import ctypes
import os
import re

from dqn_mgsc_zoo_b200 import _lib, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "dqnb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(dqnb_[a-z0-9_]+)\s*\(", text)))


def test_library_builds_and_exports_every_declared_symbol():
    build.build()
    lib = ctypes.CDLL(_lib.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 15
    for name in names:
        assert hasattr(lib, name), "libdqnb200.so does not export %s" % name


def test_ctypes_table_matches_header():
    names = set(declared_symbols()) - {"dqnb_version", "dqnb_last_error"}
    assert names == set(_lib.SIGNATURES), names ^ set(_lib.SIGNATURES)


def test_load_and_version():
    assert _lib.load().dqnb_version() >= 100
    assert _lib.load().dqnb_last_error() is not None
