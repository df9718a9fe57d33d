"""ovl_filter_rows_static -- the row filter for hosts that are not allowed to allocate.

Some programs that load libphasm_ovl.so cannot let it call cudaMalloc while they run: their own
memory pool already owns the GPU, or they sit in a latency-bound loop where the device-wide stall
of a cudaMalloc / cudaFree is unacceptable, or their memory budget was fixed when they started.
For those the library has a second filter entry point whose working memory is a fixed array inside
the library image (6 MiB per GPU, see include/phasm_ovl.h); the caller passes device pointers for
the rows, the entry lengths and the results, and the call allocates nothing.

This file checks that entry point, through the C ABI, against the allocating one
(ovl_filter_rows) on the same rows: kept rows and counters must be IDENTICAL (integer work, no
tolerance).  The GPU test issues two calls back to back on one stream -- different row sets,
different options -- before looking at either result, because the scratch array is shared and keeps
the previous call's contents: the second call must not see the first one's marks, flags or counts.

The file is named test_zz_* so that pytest collects it last: `pytest -x` stops at the first failure
and a failure here must not hide the parity tests proper.  It has no dependence on what ran before
it (fresh ExactOverlapper, own device buffers), so it passes alone in a fresh process as well.
"""
import ctypes

import numpy as np
import pytest

OPTION_SETS = (
    dict(contained_reads=True, max_overhang_abs=1000, max_overhang_rel=0.8, collapse_pairs=False),  # the CLI defaults
    dict(contained_reads=False, max_overhang_abs=None, collapse_pairs=True),
    dict(contained_reads=True, min_read_length=400, min_overlap_length=250, collapse_pairs=True),
    dict(contained_reads=False, max_overhang_abs=0, max_overhang_rel=0.0),
    dict(contained_reads=False, max_overhang_abs=-1),
)


def make_opts(contained_reads=True, min_read_length=0, min_overlap_length=0, max_overhang_abs=1000,
              max_overhang_rel=0.8, collapse_pairs=False):
    """The same mapping ExactOverlapper.filter_rows applies to its keyword arguments."""
    from phasm_b200._capi import OvlFilterOpts
    o = OvlFilterOpts()
    o.contained_reads = int(bool(contained_reads))
    o.min_read_length = int(min_read_length)
    o.min_overlap_length = int(min_overlap_length)
    o.max_overhang = 0 if max_overhang_abs is None else 1
    o.max_overhang_abs = 0 if max_overhang_abs is None else int(max_overhang_abs)
    o.max_overhang_rel = float(max_overhang_rel)
    o.collapse_pairs = int(bool(collapse_pairs))
    return o


class StaticCall:
    """Device buffers (the caller's job) + one enqueue of ovl_filter_rows_static on torch's current stream."""

    def __init__(self, lib, rows: np.ndarray, lens: np.ndarray, opt: dict):
        import torch
        self.torch = torch
        dev = torch.device("cuda", 0)
        rows = np.ascontiguousarray(rows, dtype=np.uint32).reshape(-1, 5)
        self.n = len(rows)
        self.d_rows = torch.from_numpy(rows.view(np.int32).copy()).to(dev)
        self.d_len = torch.from_numpy(np.ascontiguousarray(lens, dtype=np.uint32).view(np.int32).copy()).to(dev)
        self.d_out = torch.full((max(self.n, 1), 5), -1, dtype=torch.int32, device=dev)  # garbage the call must overwrite
        self.d_counts = torch.full((4,), -1, dtype=torch.int64, device=dev)
        self.err = ctypes.create_string_buffer(256)
        self.rc = lib.ovl_filter_rows_static(
            0, ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream),
            ctypes.c_void_p(self.d_rows.data_ptr() if self.n else 0), self.n,
            ctypes.c_void_p(self.d_len.data_ptr()), len(lens), ctypes.byref(make_opts(**opt)),
            ctypes.c_void_p(self.d_out.data_ptr()), ctypes.c_void_p(self.d_counts.data_ptr()),
            self.err, len(self.err))

    def result(self):
        assert self.rc == 0, self.err.value.decode()
        self.torch.cuda.synchronize()
        counts = [int(x) for x in self.d_counts.cpu().numpy()]
        kept = self.d_out.cpu().numpy().view(np.uint32)[:counts[0]]
        return kept, dict(rows_kept=counts[0], rows_contained=counts[1], entries_removed=counts[2], rows_bad=counts[3])


def job(ov, reads, min_length):
    ov.clear()
    for name, seq in reads:
        ov.add_read(name, seq)
    rows = ov.overlaps_arrays(min_length)
    return rows, ov.lengths()


def check(call: StaticCall, want_rows, want_st, what):
    got, st = call.result()
    assert st["rows_bad"] == 0, what
    assert st["rows_kept"] == want_st["rows_kept"] == len(want_rows), what
    assert np.array_equal(got, want_rows.reshape(-1, 5)), what
    assert st["rows_contained"] == want_st["rows_contained"], what
    assert st["entries_removed"] == want_st["entries_removed"], what


@pytest.mark.gpu
def test_static_filter_equals_allocating_filter_on_two_calls_in_a_row():
    import helpers
    from phasm_b200 import _capi, synth
    from phasm_b200.overlapper import ExactOverlapper
    lib = _capi.load_library()
    ov = ExactOverlapper()
    try:
        # job A: many contained reads (the first-mark table gets busy); job B: a different, smaller table
        rows_a, lens_a = job(ov, synth.edge_mix(5, 200), 200)
        want_a = [ov.filter_rows(**o) for o in OPTION_SETS]
        reads_b, l_b = helpers.mid_input("cfg2_x0.015")
        rows_b, lens_b = job(ov, reads_b, l_b)
        want_b = [ov.filter_rows(**o) for o in OPTION_SETS]
        assert len(rows_a) > 100 and len(rows_b) > 100 and len(lens_a) != len(lens_b)
        assert any(st["entries_removed"] for _, st in want_a)

        # two calls in a row, nothing read back in between: B must not see what A left in the scratch
        first = StaticCall(lib, rows_a, lens_a, OPTION_SETS[0])
        second = StaticCall(lib, rows_b, lens_b, OPTION_SETS[2])
        check(first, *want_a[0], "first call")
        check(second, *want_b[2], "second call")

        # ... and every option set on both jobs, alternating the jobs so that each call follows a foreign one
        for i, opt in enumerate(OPTION_SETS):
            ca = StaticCall(lib, rows_a, lens_a, opt)
            cb = StaticCall(lib, rows_b, lens_b, opt)
            check(ca, *want_a[i], ("A", opt))
            check(cb, *want_b[i], ("B", opt))

        # the same call twice gives the same answer (nothing accumulates in the scratch)
        again = [StaticCall(lib, rows_a, lens_a, OPTION_SETS[0]) for _ in range(2)]
        for c in again:
            check(c, *want_a[0], "repeat")

        # no rows: the counters still come back, all zero
        empty = StaticCall(lib, rows_a[:0], lens_a, OPTION_SETS[0])
        got, st = empty.result()
        assert len(got) == 0 and st == dict(rows_kept=0, rows_contained=0, entries_removed=0, rows_bad=0)

        # a row that names an entry outside the table is dropped and counted, not followed
        bad = rows_a.copy()
        bad[3, 1] = len(lens_a) + 7
        c = StaticCall(lib, bad, lens_a, dict(contained_reads=False, max_overhang_abs=None))
        got, st = c.result()
        assert st["rows_bad"] == 1 and st["rows_kept"] == len(bad) - 1
        assert np.array_equal(got, np.delete(bad, 3, axis=0))
    finally:
        ov.close()


def test_static_filter_rejects_jobs_beyond_its_fixed_scratch():
    """Argument and size checks come before any CUDA call, so they can be exercised without a GPU."""
    from phasm_b200 import _capi
    lib = _capi.load_library()
    err = ctypes.create_string_buffer(256)
    dummy = ctypes.c_void_p(256)  # never dereferenced: the call must fail first
    o = make_opts()
    rc = lib.ovl_filter_rows_static(0, None, dummy, _capi.OVL_STATIC_MAX_ROWS + 1, dummy, 10, ctypes.byref(o),
                                    dummy, dummy, err, len(err))
    assert rc == _capi.OVL_E_LIMIT and b"OVL_STATIC_MAX_ROWS" in err.value
    rc = lib.ovl_filter_rows_static(0, None, dummy, 10, dummy, _capi.OVL_STATIC_MAX_ENTRIES + 1, ctypes.byref(o),
                                    dummy, dummy, err, len(err))
    assert rc == _capi.OVL_E_LIMIT and b"OVL_STATIC_MAX_ENTRIES" in err.value
    rc = lib.ovl_filter_rows_static(0, None, dummy, 10, dummy, 10, None, dummy, dummy, err, len(err))
    assert rc == _capi.OVL_E_ARG and b"NULL" in err.value
    rc = lib.ovl_filter_rows_static(0, None, dummy, 10, dummy, 10, ctypes.byref(o), dummy, None, None, 0)
    assert rc == _capi.OVL_E_ARG   # no message buffer is fine too
