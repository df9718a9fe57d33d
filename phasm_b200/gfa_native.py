"""Native GFA2 ingest (parsing part): ``ovl_gfa_read`` through ctypes.

Replaces the two Python passes of the reference's ``layout`` stage over the GFA2 file
(``phasm/cli/assembler.py:55-60,93-95`` over ``phasm/io/gfa.py:34-47,65-107``) with one native
parse into arrays.  Host only: works without a GPU.
"""
import ctypes
from ctypes import byref, c_size_t, c_uint32, c_uint64, c_void_p
from typing import List, Tuple

import numpy as np

from . import _capi

EDGE_DTYPE = np.dtype([("a", "<u4"), ("b", "<u4"), ("astart", "<u4"), ("aend", "<u4"),
                       ("bstart", "<u4"), ("bend", "<u4"), ("a_strand", "S1"), ("b_strand", "S1"),
                       ("reserved", "V2")])
assert EDGE_DTYPE.itemsize == ctypes.sizeof(_capi.OvlGfaEdge)


def read_gfa(path: str) -> Tuple[List[str], np.ndarray, np.ndarray]:
    """-> (segment names, segment lengths uint32[S], edges EDGE_DTYPE[E] on segment indices, file order).

    Raises ``ValueError`` for a malformed line or an edge naming an unknown segment (the reference
    raises ``IndexError`` / ``ValueError`` / ``KeyError`` there)."""
    lib = _capi.load_library()
    g = c_void_p()
    rc = lib.ovl_gfa_read(str(path).encode(), byref(g))
    if rc != 0:
        _capi.raise_for(rc, (lib.ovl_last_error(None) or b"").decode("utf-8", "replace"))
    try:
        ns, ne = c_uint32(), c_uint64()
        lib.ovl_gfa_counts(g, byref(ns), byref(ne))
        names, lengths = [], np.zeros(ns.value, dtype=np.uint32)
        p, n, L = c_void_p(), c_size_t(), c_uint32()
        for i in range(ns.value):
            lib.ovl_gfa_segment(g, i, byref(p), byref(n), byref(L))
            names.append(ctypes.string_at(p.value, n.value).decode("utf-8") if n.value else "")
            lengths[i] = L.value
        ep, en = c_void_p(), c_uint64()
        lib.ovl_gfa_edges(g, byref(ep), byref(en))
        if en.value:
            buf = ctypes.string_at(ep.value, en.value * EDGE_DTYPE.itemsize)
            edges = np.frombuffer(buf, dtype=EDGE_DTYPE).copy()
        else:
            edges = np.zeros(0, dtype=EDGE_DTYPE)
        return names, lengths, edges
    finally:
        lib.ovl_gfa_free(g)
