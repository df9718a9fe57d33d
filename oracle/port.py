"""ctypes front-end of the T2 oracle (``oracle/overlap_oracle.c``).

TEST INFRASTRUCTURE ONLY: importable from ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s CPU-baseline legs; never from ``phasm_b200/``.

``oracle_overlaps(buf, offsets, min_length)`` returns ``(rows, stats)`` where ``rows`` is a
``(n, 5)`` uint32 array of ``(a, b, astart, aend, bend)`` on entry indices (bstart is always
0, /root/reference/src/overlapper.cpp:81,109), sorted lexicographically, and ``stats`` holds
the data-determined counts used for the algorithmic-byte roofline.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import Dict, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboverlap_oracle.so")


class _Stats(ctypes.Structure):
    _fields_ = [(n, ctypes.c_uint64) for n in
                ("anchors", "positions", "candidates", "compared_bases", "operand_bytes",
                 "hits", "rows")]


_lib = None


def build(force: bool = False) -> str:
    """Compile the C restatement (gcc only; no reference files needed)."""
    src = os.path.join(_HERE, "overlap_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "port"], stdout=subprocess.DEVNULL)
    return _SO


def _load():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        lib = ctypes.CDLL(_SO)
        lib.ovo_overlaps.restype = ctypes.c_int
        lib.ovo_overlaps.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32,
                                     ctypes.c_uint32, ctypes.c_int,
                                     ctypes.POINTER(ctypes.c_void_p),
                                     ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(_Stats)]
        lib.ovo_free.restype = None
        lib.ovo_free.argtypes = [ctypes.c_void_p]
        _lib = lib
    return _lib


def oracle_overlaps(buf: np.ndarray, offsets: np.ndarray, min_length: int,
                    nthreads: int = 0) -> Tuple[np.ndarray, Dict[str, int]]:
    """buf: uint8 concatenated entries; offsets: uint64[M+1] entry boundaries."""
    lib = _load()
    buf = np.ascontiguousarray(buf, dtype=np.uint8)
    offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
    m = len(offsets) - 1
    rows_p = ctypes.c_void_p()
    n = ctypes.c_uint64()
    st = _Stats()
    rc = lib.ovo_overlaps(buf.ctypes.data, offsets.ctypes.data, m, int(min_length), nthreads,
                          ctypes.byref(rows_p), ctypes.byref(n), ctypes.byref(st))
    if rc != 0:
        raise MemoryError("oracle: allocation failed")
    try:
        if n.value:
            arr = np.ctypeslib.as_array(
                ctypes.cast(rows_p, ctypes.POINTER(ctypes.c_uint32)), shape=(n.value, 5)).copy()
        else:
            arr = np.zeros((0, 5), dtype=np.uint32)
    finally:
        lib.ovo_free(rows_p)
    stats = {f: int(getattr(st, f)) for f, _ in _Stats._fields_}
    return arr, stats


def entries_to_buffer(entries) -> Tuple[np.ndarray, np.ndarray]:
    """[(id, seq str|bytes)] -> (uint8 buffer, uint64 offsets)."""
    bs = [s.encode("utf-8") if isinstance(s, str) else bytes(s) for _, s in entries]
    offsets = np.zeros(len(bs) + 1, dtype=np.uint64)
    if bs:
        offsets[1:] = np.cumsum([len(b) for b in bs], dtype=np.uint64)
    buf = np.frombuffer(b"".join(bs), dtype=np.uint8) if bs else np.zeros(0, dtype=np.uint8)
    return buf, offsets


def rows_to_tuples(rows: np.ndarray, ids) -> list:
    """(n,5) index rows -> the reference's tuple shape (overlapper.h:17), sorted."""
    out = [(ids[a], ids[b], int(s), int(e), 0, int(be)) for a, b, s, e, be in rows.tolist()]
    out.sort()
    return out
