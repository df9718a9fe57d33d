"""T0 oracle: the UNMODIFIED reference overlapper, compiled from its own sources.

TEST INFRASTRUCTURE ONLY (tests/, smoke(), bench.py CPU-baseline legs).

``oracle/Makefile`` (target ``ref``) compiles /root/reference/src/{overlapper,phasm}.cpp with
the vendored SeqAn headers into ``oracle/_ref/overlapper<EXT_SUFFIX>``; this module only loads
that prebuilt extension (the GPU box has no /root/reference, the .so travels with the repo).
"""
from __future__ import annotations

import importlib.util
import glob
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_mod = None


def available() -> bool:
    return bool(glob.glob(os.path.join(_HERE, "_ref", "overlapper*.so")))


def build() -> None:
    """Build oracle/_ref from /root/reference (only possible where the reference is mounted)."""
    if available():
        return
    if not os.path.isdir("/root/reference/src"):
        raise RuntimeError("reference sources not mounted; oracle/_ref cannot be built here")
    subprocess.check_call(["make", "-C", _HERE, "ref"], stdout=subprocess.DEVNULL)


def load():
    """Return the compiled reference module (``ExactOverlapper`` class inside)."""
    global _mod
    if _mod is None:
        sos = glob.glob(os.path.join(_HERE, "_ref", "overlapper*.so"))
        if not sos:
            raise ImportError("oracle/_ref not built (run `make -C oracle ref` where "
                              "/root/reference is mounted)")
        spec = importlib.util.spec_from_file_location("overlapper", sos[0])
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        _mod = mod
    return _mod


def ref_overlaps(entries, min_length: int) -> list:
    """Drive the reference exactly as assembler.py:39-42 does; returns sorted tuples.

    The reference segfaults when every entry is empty; callers avoid that.
    """
    ov = load().ExactOverlapper()
    for eid, seq in entries:
        ov.add_sequence(eid, seq)
    return sorted(ov.overlaps(int(min_length)))
