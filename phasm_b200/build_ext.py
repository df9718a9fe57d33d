"""Build the drop-in extension module ``phasm/overlapper<EXT_SUFFIX>`` (pybind11 over the C ABI).

Replaces the reference's ``setup.py:44-70`` extension ``phasm.overlapper``: same module name,
same class, but the object code is ``csrc/pyext.cpp`` linked against the in-tree
``phasm_b200/_lib/libphasm_ovl.so`` (found at run time through an ``$ORIGIN`` rpath, so the
pair travels together).
"""
from __future__ import annotations

import os
import subprocess
import sysconfig

_HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(_HERE)
SRC = os.path.join(_HERE, "csrc", "pyext.cpp")
HDR = os.path.join(ROOT, "include", "phasm_ovl.h")
LIBDIR = os.path.join(_HERE, "_lib")


def target() -> str:
    return os.path.join(ROOT, "phasm", "overlapper" + sysconfig.get_config_var("EXT_SUFFIX"))


def build(force: bool = False) -> str:
    import pybind11
    out = target()
    lib = os.path.join(LIBDIR, "libphasm_ovl.so")
    if not os.path.exists(lib):
        raise RuntimeError("build libphasm_ovl.so first (__graft_entry__.build_library)")
    if (not force and os.path.exists(out)
            and all(os.path.getmtime(s) <= os.path.getmtime(out) for s in (SRC, HDR, lib))):
        return out
    cxx = os.environ.get("CXX", "g++")
    cmd = [cxx, "-O2", "-std=c++17", "-fPIC", "-shared", "-fvisibility=hidden", "-Wall",
           "-I" + sysconfig.get_paths()["include"], "-I" + pybind11.get_include(),
           SRC, "-o", out, "-L" + LIBDIR, "-lphasm_ovl",
           "-Wl,-rpath,$ORIGIN/../phasm_b200/_lib"]
    subprocess.check_call(cmd)
    return out
