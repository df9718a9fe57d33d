"""Host-side mirror of the reference's pybind11 module ``phasm.overlapper``.

Reference interface (/root/reference/src/phasm.cpp:8-18):

    ExactOverlapper()                       no-arg constructor         (phasm.cpp:13)
    .add_sequence(id, seq) -> None          positional, copies both    (phasm.cpp:14)
    .overlaps(min_length) -> list[tuple[str, str, int, int, int, int]] (phasm.cpp:15)

Same names, argument meaning and error behaviour; the work happens in hand-written sm_100a
CUDA behind the C ABI of ``include/phasm_ovl.h`` (``libphasm_ovl.so``).  There is no CPU
path: constructing an ``ExactOverlapper`` without a usable B200 raises.

Additive extras (not in the reference; defaults keep reference behaviour):
``add_read`` (both strands, reverse complement made on the device), ``add_reads_bulk``,
``overlaps_arrays`` (index rows as a numpy array instead of 10^7 Python tuples), ``stats``.
"""
from __future__ import annotations

import ctypes
import numbers
from ctypes import byref, c_size_t, c_uint32, c_uint64, c_void_p
from typing import List, Optional, Sequence, Tuple

import numpy as np

from . import _capi
from ._capi import OvlRow, OvlStats

__doc__module__ = "Optimised C++ all-vs-all exact overlap finder for PHASM"  # phasm.cpp:9-10

Row = Tuple[str, str, int, int, int, int]

# one edge of the assembly graph (ovl_edge in include/phasm_ovl.h)
EDGE_DTYPE = np.dtype([("u", "<u4"), ("v", "<u4"), ("edge_len", "<i4"), ("overlap_len", "<u4"), ("first_seq", "<u8")])
assert EDGE_DTYPE.itemsize == ctypes.sizeof(_capi.OvlEdge)


def _as_bytes(x, what: str) -> bytes:
    if isinstance(x, str):
        return x.encode("utf-8")
    if isinstance(x, (bytes, bytearray, memoryview)):
        return bytes(x)
    raise TypeError("%s must be str or bytes, not %s" % (what, type(x).__name__))


def _as_unsigned(x, what: str) -> int:
    # pybind11 converts to `unsigned int`: bool/int accepted, float and negatives rejected
    if isinstance(x, bool) or not isinstance(x, numbers.Integral):
        if isinstance(x, bool):
            return int(x)
        raise TypeError("%s: incompatible function arguments (expected unsigned int, got %s)"
                        % (what, type(x).__name__))
    x = int(x)
    if x < 0 or x > 0xFFFFFFFF:
        raise TypeError("%s: incompatible function arguments (value %d does not fit unsigned int)"
                        % (what, x))
    return x


class ExactOverlapper:
    """All-vs-all exact suffix-prefix overlap finder (drop-in for phasm.overlapper.ExactOverlapper)."""

    def __init__(self, device: int = 0):
        self._lib = _capi.load_library()
        self._ctx = c_void_p()
        rc = self._lib.ovl_create(int(device), byref(self._ctx))
        if rc != _capi.OVL_OK:
            msg = (self._lib.ovl_last_error(None) or b"").decode("utf-8", "replace")
            self._ctx = c_void_p()
            _capi.raise_for(rc, "ExactOverlapper: " + msg)

    # ---- plumbing -----------------------------------------------------------------------
    def _check(self, rc: int):
        if rc != _capi.OVL_OK:
            msg = (self._lib.ovl_last_error(self._ctx) or b"").decode("utf-8", "replace")
            _capi.raise_for(rc, msg)

    def close(self):
        if getattr(self, "_ctx", None) is not None and self._ctx.value:
            self._lib.ovl_destroy(self._ctx)
            self._ctx = c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- reference API ------------------------------------------------------------------
    def add_sequence(self, id, seq, /) -> None:
        """ExactOverlapper::addSequence (src/overlapper.cpp:22-26): append a copy of (id, seq)."""
        bid, bseq = _as_bytes(id, "id"), _as_bytes(seq, "seq")
        self._check(self._lib.ovl_add_sequence(self._ctx, bid, len(bid), bseq, len(bseq)))

    def overlaps(self, min_length, /) -> List[Row]:
        """ExactOverlapper::overlaps (src/overlapper.cpp:28-150).

        Returns ``[(a_id, b_id, astart, aend, bstart, bend)]`` with bstart always 0.  The
        reference's row order is unspecified; here rows are sorted by entry index, position.
        """
        rows = self.overlaps_arrays(min_length)
        ids = self.ids()
        a = rows[:, 0].tolist()
        b = rows[:, 1].tolist()
        s = rows[:, 2].tolist()
        e = rows[:, 3].tolist()
        be = rows[:, 4].tolist()
        return [(ids[a[i]], ids[b[i]], s[i], e[i], 0, be[i]) for i in range(len(a))]

    # ---- additive extras ------------------------------------------------------------------
    def add_read(self, name, seq) -> None:
        """What phasm/cli/assembler.py:39-40 does per FASTA record: entries name+"+" (seq) and
        name+"-" (reverse complement, computed on the device)."""
        bn, bs = _as_bytes(name, "name"), _as_bytes(seq, "seq")
        self._check(self._lib.ovl_add_read(self._ctx, bn, len(bn), bs, len(bs)))

    def add_reads_bulk(self, buf, offsets, names: Optional[Sequence[str]] = None,
                       both_strands: bool = True) -> None:
        """Add ``len(offsets)-1`` records at once: record i is ``buf[offsets[i]:offsets[i+1]]``."""
        buf = np.ascontiguousarray(buf, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        n = len(offsets) - 1
        if n <= 0:
            return
        if names is not None:
            enc = [_as_bytes(x, "name") for x in names]
            if len(enc) != n:
                raise ValueError("names must have one entry per record")
            noff = np.zeros(n + 1, dtype=np.uint64)
            noff[1:] = np.cumsum([len(x) for x in enc], dtype=np.uint64)
            blob = b"".join(enc)
            nptr = ctypes.cast(ctypes.c_char_p(blob), c_void_p)
            self._check(self._lib.ovl_add_bulk(self._ctx, buf.ctypes.data, offsets.ctypes.data, n,
                                               nptr, noff.ctypes.data, int(both_strands)))
        else:
            self._check(self._lib.ovl_add_bulk(self._ctx, buf.ctypes.data, offsets.ctypes.data, n,
                                               None, None, int(both_strands)))

    def add_fasta(self, path, both_strands: bool = True) -> int:
        """Native FASTA ingest (``ovl_add_fasta``): what the dinopy loop of
        phasm/cli/assembler.py:32-40 does, without Python objects per read.  Returns the number
        of records."""
        import os
        n = c_uint32()
        self._check(self._lib.ovl_add_fasta(self._ctx, os.fsencode(path), int(both_strands), byref(n)))
        return n.value

    def add_fasta_shard(self, path, rank: int, world: int, both_strands: bool = True):
        """Multi-GPU ingest (``ovl_add_fasta_shard``): stage this rank's record range of the file,
        declare the other records by name and length.  Returns ``(n_records, lo, hi)``."""
        import os
        n, lo, hi = c_uint32(), c_uint32(), c_uint32()
        self._check(self._lib.ovl_add_fasta_shard(self._ctx, os.fsencode(path), int(both_strands), int(rank),
                                                  int(world), byref(n), byref(lo), byref(hi)))
        return n.value, lo.value, hi.value

    def write_gfa(self, out, header: bool = True, segments: bool = True, edges: bool = True) -> int:
        """Native GFA2 writer (``ovl_write_gfa``): header / S-lines / E-lines of the last
        ``overlaps*`` call, byte-identical to phasm/cli/assembler.py:30,38,46-48.  ``out`` is a
        file descriptor or a file object with ``fileno()`` (flushed first).  Returns bytes written."""
        if hasattr(out, "fileno"):
            out.flush()
            fd = out.fileno()
        else:
            fd = int(out)
        nb = c_uint64()
        what = (1 if header else 0) | (2 if segments else 0) | (4 if edges else 0)
        self._check(self._lib.ovl_write_gfa(self._ctx, fd, what, byref(nb)))
        return nb.value

    def add_remote(self, lengths, both_strands: bool = True) -> None:
        """Multi-GPU: declare records held by another rank (lengths only)."""
        lengths = np.ascontiguousarray(lengths, dtype=np.uint32)
        if len(lengths):
            self._check(self._lib.ovl_add_remote(self._ctx, lengths.ctypes.data, len(lengths),
                                                 int(both_strands)))

    def clear(self) -> None:
        self._check(self._lib.ovl_clear(self._ctx))

    def __len__(self) -> int:
        m = c_uint32()
        self._check(self._lib.ovl_entry_count(self._ctx, byref(m)))
        return m.value

    def ids(self) -> List[str]:
        out = []
        p, n = c_void_p(), c_size_t()
        for i in range(len(self)):
            self._check(self._lib.ovl_entry_id(self._ctx, i, byref(p), byref(n)))
            out.append(ctypes.string_at(p.value, n.value).decode("utf-8") if n.value else "")
        return out

    def lengths(self) -> np.ndarray:
        m = len(self)
        out = np.empty(m, dtype=np.uint32)
        v = c_uint32()
        for i in range(m):
            self._check(self._lib.ovl_entry_length(self._ctx, i, byref(v)))
            out[i] = v.value
        return out

    @staticmethod
    def _rows_array(p: c_void_p, n: int, copy: bool = True) -> np.ndarray:
        if not n:
            return np.zeros((0, 5), dtype=np.uint32)
        arr = np.ctypeslib.as_array(ctypes.cast(p, ctypes.POINTER(c_uint32)), shape=(n, 5))
        if copy:
            return arr.copy()
        arr.flags.writeable = False
        return arr

    def overlaps_arrays(self, min_length, copy: bool = True) -> np.ndarray:
        """Rows as a ``(n, 5)`` uint32 array ``(a, b, astart, aend, bend)`` on entry indices.

        ``copy=False`` returns a read-only view of the library's pinned result buffer (what the C
        ABI hands out: ``ovl_overlaps`` returns a pointer it owns); it is valid until the next
        call on this object."""
        l = _as_unsigned(min_length, "overlaps")
        p, n = c_void_p(), c_uint64()
        self._check(self._lib.ovl_overlaps(self._ctx, l, byref(p), byref(n)))
        return self._rows_array(p, n.value, copy)

    # split form (device-resident timing)
    def upload(self) -> None:
        self._check(self._lib.ovl_upload(self._ctx))

    def overlaps_device(self, min_length) -> int:
        """One pass over HBM-resident inputs; the rows stay in HBM.  Multi-GPU: collective; every
        rank ends with the rows of its A range in its own HBM, rank 0 returns the job's row count
        (the others 0) and ``fetch_rows`` on rank 0 pulls all of them over NVLink."""
        n = c_uint64()
        self._check(self._lib.ovl_overlaps_device(self._ctx, _as_unsigned(min_length, "overlaps"),
                                                  byref(n)))
        return n.value

    def fetch_rows(self, copy: bool = True) -> np.ndarray:
        p, n = c_void_p(), c_uint64()
        self._check(self._lib.ovl_fetch_rows(self._ctx, byref(p), byref(n)))
        return self._rows_array(p, n.value, copy)

    def fetch_local_rows(self, copy: bool = True) -> np.ndarray:
        """Multi-GPU: this rank's OWN rows of the last pass (``stats()["rows_local"]`` of them),
        HBM -> this rank's pinned host buffer over its own PCIe link."""
        p, n = c_void_p(), c_uint64()
        self._check(self._lib.ovl_fetch_local_rows(self._ctx, byref(p), byref(n)))
        return self._rows_array(p, n.value, copy)

    def overlaps_local_arrays(self, min_length, copy: bool = True) -> np.ndarray:
        """Multi-GPU, collective: host buffers in on every rank, the pass, this rank's own rows out
        (``ovl_overlaps_local``).  The ranks' arrays concatenated in rank order are the canonical
        row list; rank 0 also holds all rows in HBM (``fetch_rows``)."""
        l = _as_unsigned(min_length, "overlaps")
        p, n = c_void_p(), c_uint64()
        self._check(self._lib.ovl_overlaps_local(self._ctx, l, byref(p), byref(n)))
        return self._rows_array(p, n.value, copy)

    def filter_rows(self, contained_reads: bool = True, min_read_length: int = 0,
                    min_overlap_length: int = 0, max_overhang_abs=1000, max_overhang_rel=0.8,
                    collapse_pairs: bool = False, copy: bool = True):
        """Optional device post-pass over the rows of the last ``overlaps*`` call: the alignment
        filters of the reference's ``layout`` stage, in its order and with its streaming semantics
        (``phasm/cli/assembler.py:76-86``: ``ContainedReads()``, ``MinReadLength``,
        ``MinOverlapLength``, ``MaxOverhang(abs, rel)``; defaults as the reference CLI,
        ``assembler.py:469-490``).  ``max_overhang_abs=None`` switches MaxOverhang off.
        Returns ``(rows, stats)``: the kept rows in canonical order and a dict of counters."""
        from ._capi import OvlFilterOpts, OvlFilterStats
        o = OvlFilterOpts()
        o.contained_reads = int(bool(contained_reads))
        o.min_read_length = int(min_read_length)
        o.min_overlap_length = int(min_overlap_length)
        o.max_overhang = 0 if max_overhang_abs is None else 1
        o.max_overhang_abs = 0 if max_overhang_abs is None else int(max_overhang_abs)
        o.max_overhang_rel = float(max_overhang_rel)
        o.collapse_pairs = int(bool(collapse_pairs))
        st = OvlFilterStats()
        p, n = c_void_p(), c_uint64()
        self._check(self._lib.ovl_filter_rows(self._ctx, byref(o), byref(p), byref(n), byref(st)))
        return self._rows_array(p, n.value, copy), st.as_dict()

    def build_graph(self, filtered: bool = True):
        """Optional device post-pass: the edge list of the assembly graph the reference's ``layout``
        stage builds from the alignments (``build_assembly_graph``, ``phasm/assembly_graph.py:136-179``)
        -- over the rows kept by the last ``filter_rows`` call (``filtered=True``, what
        ``assembler.py:93-100`` feeds it) or over all rows of the last ``overlaps*`` call.
        Returns ``(edges, stats)``: a structured array (``EDGE_DTYPE``: u, v, edge_len, overlap_len,
        first_seq) sorted by (u, v) on entry indices, and a dict with networkx's node / edge counts."""
        from ._capi import OvlGraphStats
        st = OvlGraphStats()
        p, n = c_void_p(), c_uint64()
        self._check(self._lib.ovl_build_graph(self._ctx, int(bool(filtered)), byref(p), byref(n), byref(st)))
        if n.value:
            buf = ctypes.string_at(p.value, n.value * EDGE_DTYPE.itemsize)
            edges = np.frombuffer(buf, dtype=EDGE_DTYPE).copy()
        else:
            edges = np.zeros(0, dtype=EDGE_DTYPE)
        return edges, st.as_dict()

    def set_collect_stats(self, on: bool) -> None:
        self._check(self._lib.ovl_set_collect_stats(self._ctx, int(bool(on))))

    def set_option(self, name: str, value: int) -> None:
        """Tuning knob of the C ABI (``ovl_set_option``), e.g. ``anchor_stride``."""
        self._check(self._lib.ovl_set_option(self._ctx, name.encode(), int(value)))

    def stats(self) -> dict:
        st = OvlStats()
        self._check(self._lib.ovl_get_stats(self._ctx, byref(st)))
        return st.as_dict()

    def timer_start(self) -> None:
        """Record a CUDA event on the stream this object's kernels are launched on."""
        self._check(self._lib.ovl_timer_start(self._ctx))

    def timer_stop(self) -> float:
        """Record a second event, wait for it; milliseconds since ``timer_start``."""
        ms = ctypes.c_float()
        self._check(self._lib.ovl_timer_stop(self._ctx, byref(ms)))
        return ms.value

    def debug_packed(self, idx: int) -> np.ndarray:
        n = c_size_t()
        cap = (int(self.lengths()[idx]) + 63) // 64 * 16
        out = np.zeros(max(cap, 1), dtype=np.uint8)
        self._check(self._lib.ovl_debug_packed(self._ctx, idx, out.ctypes.data, cap, byref(n)))
        return out[:n.value]

    # ---- multi-GPU (one process per GPU) ---------------------------------------------------
    @staticmethod
    def comm_unique_id() -> bytes:
        lib = _capi.load_library()
        buf = (ctypes.c_uint8 * 128)()
        rc = lib.ovl_comm_unique_id(buf)
        if rc != _capi.OVL_OK:
            _capi.raise_for(rc, (lib.ovl_last_error(None) or b"").decode("utf-8", "replace"))
        return bytes(buf)

    def comm_init(self, uid: bytes, rank: int, nranks: int) -> None:
        buf = (ctypes.c_uint8 * 128).from_buffer_copy(uid)
        self._check(self._lib.ovl_comm_init(self._ctx, buf, int(rank), int(nranks)))

    def comm_sync_ranges(self) -> None:
        self._check(self._lib.ovl_comm_sync_ranges(self._ctx))

    def comm_gather_rows_device(self) -> int:
        n = c_uint64()
        self._check(self._lib.ovl_comm_gather_rows_device(self._ctx, byref(n)))
        return n.value

    def comm_gather_rows(self, copy: bool = True) -> np.ndarray:
        p, n = c_void_p(), c_uint64()
        self._check(self._lib.ovl_comm_gather_rows(self._ctx, byref(p), byref(n)))
        return self._rows_array(p, n.value, copy)
