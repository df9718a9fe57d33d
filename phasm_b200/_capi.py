"""ctypes binding of ``libphasm_ovl.so`` (the C ABI declared in ``include/phasm_ovl.h``).

No torch, no numpy requirement on the call path (numpy is used only to hand out row arrays).
There is no CPU fallback: if the library is missing or no CUDA device is usable, the calls
raise.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import (POINTER, Structure, byref, c_char_p, c_float, c_int, c_size_t, c_uint8,
                    c_uint32, c_uint64, c_void_p)

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_lib", "libphasm_ovl.so")

OVL_OK = 0
OVL_E_CUDA, OVL_E_NOMEM, OVL_E_ALPHABET, OVL_E_ARG, OVL_E_LIMIT, OVL_E_NCCL = -1, -2, -3, -4, -5, -6


class OvlRow(Structure):
    _fields_ = [(n, c_uint32) for n in ("a", "b", "astart", "aend", "bend")]


class OvlStats(Structure):
    _fields_ = ([(n, c_uint64) for n in
                 ("entries", "bases", "ascii_bytes", "granules", "anchors",
                  "candidates", "hits", "rows", "compared_bases", "operand_bytes",
                  "scan_positions")] +
                [(n, c_uint32) for n in ("kmer", "anchor_stride", "launches", "lib_launches",
                                         "retries", "mirror")] +
                [("verified", c_uint64), ("hits_sent", c_uint64), ("hits_received", c_uint64)] +
                [(n, c_float) for n in ("ms_h2d", "ms_pack", "ms_anchor", "ms_scan", "ms_sort",
                                        "ms_verify", "ms_emit", "ms_d2h", "ms_device", "ms_total")] +
                [("big_buckets", c_uint32), ("reserved0", c_uint32),
                 ("scan_survivors", c_uint64), ("rows_local", c_uint64),
                 ("ms_gather", c_float), ("ms_exposed", c_float)])

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class OvlGfaEdge(Structure):
    _fields_ = [(n, c_uint32) for n in ("a", "b", "astart", "aend", "bstart", "bend")] + \
               [("a_strand", ctypes.c_uint8), ("b_strand", ctypes.c_uint8), ("reserved", ctypes.c_uint8 * 2)]


class OvlFilterOpts(Structure):
    _fields_ = [("contained_reads", c_uint32), ("min_read_length", c_uint32),
                ("min_overlap_length", c_uint32), ("max_overhang", c_uint32),
                ("max_overhang_abs", ctypes.c_int64), ("max_overhang_rel", ctypes.c_double),
                ("collapse_pairs", c_uint32), ("reserved0", c_uint32)]


class OvlFilterStats(Structure):
    _fields_ = [("rows_in", c_uint64), ("rows_kept", c_uint64), ("rows_contained", c_uint64),
                ("entries_removed", c_uint64), ("ms_device", c_float), ("reserved0", c_uint32)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_ if n != "reserved0"}


class OvlStaticCounts(Structure):
    _fields_ = [(n, c_uint64) for n in ("rows_kept", "rows_contained", "entries_removed", "rows_bad")]


OVL_STATIC_MAX_ROWS = 1 << 21
OVL_STATIC_MAX_ENTRIES = 1 << 19


class OvlEdge(Structure):
    _fields_ = [("u", c_uint32), ("v", c_uint32), ("edge_len", ctypes.c_int32), ("overlap_len", c_uint32),
                ("first_seq", c_uint64)]


class OvlGraphStats(Structure):
    _fields_ = [("rows_in", c_uint64), ("rows_overlap", c_uint64), ("n_edges", c_uint64),
                ("n_nodes", c_uint64), ("ms_device", c_float), ("reserved0", c_uint32)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_ if n != "reserved0"}


# every symbol include/phasm_ovl.h declares: name -> (restype, argtypes)
_P = c_void_p
SIGNATURES = {
    "ovl_create": (c_int, [c_int, POINTER(_P)]),
    "ovl_destroy": (None, [_P]),
    "ovl_last_error": (c_char_p, [_P]),
    "ovl_add_sequence": (c_int, [_P, c_char_p, c_size_t, c_char_p, c_size_t]),
    "ovl_add_read": (c_int, [_P, c_char_p, c_size_t, c_char_p, c_size_t]),
    "ovl_add_bulk": (c_int, [_P, c_void_p, c_void_p, c_uint32, c_void_p, c_void_p, c_int]),
    "ovl_reserve": (c_int, [_P, c_uint64, c_uint32]),
    "ovl_add_remote": (c_int, [_P, c_void_p, c_uint32, c_int]),
    "ovl_clear": (c_int, [_P]),
    "ovl_entry_count": (c_int, [_P, POINTER(c_uint32)]),
    "ovl_entry_id": (c_int, [_P, c_uint32, POINTER(c_void_p), POINTER(c_size_t)]),
    "ovl_entry_length": (c_int, [_P, c_uint32, POINTER(c_uint32)]),
    "ovl_overlaps": (c_int, [_P, c_uint32, POINTER(c_void_p), POINTER(c_uint64)]),
    "ovl_upload": (c_int, [_P]),
    "ovl_overlaps_device": (c_int, [_P, c_uint32, POINTER(c_uint64)]),
    "ovl_fetch_rows": (c_int, [_P, POINTER(c_void_p), POINTER(c_uint64)]),
    "ovl_set_collect_stats": (c_int, [_P, c_int]),
    "ovl_set_option": (c_int, [_P, c_char_p, ctypes.c_int64]),
    "ovl_get_stats": (c_int, [_P, POINTER(OvlStats)]),
    "ovl_timer_start": (c_int, [_P]),
    "ovl_timer_stop": (c_int, [_P, POINTER(c_float)]),
    "ovl_add_fasta": (c_int, [_P, c_char_p, c_int, POINTER(c_uint32)]),
    "ovl_add_fasta_shard": (c_int, [_P, c_char_p, c_int, c_int, c_int, POINTER(c_uint32), POINTER(c_uint32),
                                    POINTER(c_uint32)]),
    "ovl_write_gfa": (c_int, [_P, c_int, c_int, POINTER(c_uint64)]),
    "ovl_filter_rows": (c_int, [_P, POINTER(OvlFilterOpts), POINTER(c_void_p), POINTER(c_uint64),
                                POINTER(OvlFilterStats)]),
    "ovl_filter_rows_static": (c_int, [c_int, c_void_p, c_void_p, c_uint64, c_void_p, c_uint32,
                                       POINTER(OvlFilterOpts), c_void_p, c_void_p, c_char_p, c_size_t]),
    "ovl_build_graph": (c_int, [_P, c_int, POINTER(c_void_p), POINTER(c_uint64), POINTER(OvlGraphStats)]),
    "ovl_gfa_read": (c_int, [c_char_p, POINTER(_P)]),
    "ovl_gfa_free": (None, [_P]),
    "ovl_gfa_counts": (c_int, [_P, POINTER(c_uint32), POINTER(c_uint64)]),
    "ovl_gfa_segment": (c_int, [_P, c_uint32, POINTER(c_void_p), POINTER(c_size_t), POINTER(c_uint32)]),
    "ovl_gfa_edges": (c_int, [_P, POINTER(c_void_p), POINTER(c_uint64)]),
    "ovl_comm_unique_id": (c_int, [c_void_p]),
    "ovl_comm_init": (c_int, [_P, c_void_p, c_int, c_int]),
    "ovl_comm_sync_ranges": (c_int, [_P]),
    "ovl_comm_gather_rows_device": (c_int, [_P, POINTER(c_uint64)]),
    "ovl_comm_gather_rows": (c_int, [_P, POINTER(c_void_p), POINTER(c_uint64)]),
    "ovl_fetch_local_rows": (c_int, [_P, POINTER(c_void_p), POINTER(c_uint64)]),
    "ovl_overlaps_local": (c_int, [_P, c_uint32, POINTER(c_void_p), POINTER(c_uint64)]),
    "ovl_debug_packed": (c_int, [_P, c_uint32, c_void_p, c_size_t, POINTER(c_size_t)]),
    "ovl_debug_mirror_items": (c_int, [c_int, c_void_p, c_uint32, c_uint32, c_uint32, c_uint32, c_uint32,
                                       POINTER(c_uint32), POINTER(c_int)]),
}

_lib = None


def load_library() -> ctypes.CDLL:
    """dlopen the in-tree library and declare every prototype.  Raises if it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                "%s not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU fallback)" % LIB_PATH)
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)      # AttributeError if the library does not export it
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


class OverlapperError(RuntimeError):
    pass


def raise_for(code: int, msg: str):
    """Map OVL_E_* to the exceptions pybind11 would raise for the reference's C++ errors
    (std::bad_alloc -> MemoryError, std::exception -> RuntimeError)."""
    if code == OVL_E_NOMEM:
        raise MemoryError(msg)
    if code == OVL_E_ALPHABET:
        raise ValueError(msg)
    if code == OVL_E_ARG:
        raise ValueError(msg)
    raise OverlapperError(msg)
