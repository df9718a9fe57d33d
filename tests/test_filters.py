"""f3 row filters: the CPU restatement against golden vectors made with the
reference's own filter classes, and the device post-pass (ovl_filter_rows) against both."""
import json
import os

import numpy as np
import pytest

import helpers
from helpers import sort_rows
from oracle import filters as F

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "filters.json")


@pytest.fixture(scope="module")
def gold():
    with open(GOLD) as f:
        return json.load(f)


def expected_kept(c):
    if c["kept"] is not None:
        return list(c["kept"])
    d = set(c["dropped"])
    return [i for i in range(c["n_rows"]) if i not in d]


def case_input(c):
    """-> (rows, lens, reads or None) of a chain case."""
    if c["case"] == "tiny":
        return [tuple(r) for r in c["rows"]], c["lens"], None
    reads, _ = helpers.mid_input(c["case"])
    lens = []
    for _, s in reads:
        lens += [len(s), len(s)]
    rows = [tuple(int(x) for x in r) for r in sort_rows(helpers.load_mid()[c["case"]])]
    return rows, lens, reads


def test_formulas_match_reference_local_alignment(gold):
    for a0, a1, b0, b1, la, lb, typ, olen, ohang in gold["la"]:
        assert F.classify(a0, a1, b0, b1, la, lb) == typ
        assert F.overlap_length(a0, a1, b0, b1) == olen
        assert F.overhang(a0, a1, b0, b1, la, lb) == ohang


def test_restatement_matches_reference_filter_chain(gold):
    opts = gold["option_sets"]
    n = 0
    for c in gold["chain"]:
        rows, lens, _ = case_input(c)
        kept, st = F.filter_rows(rows, lens, **opts[c["opt"]])
        assert kept == expected_kept(c), (c["case"], c["opt"])
        assert st["rows_contained"] == c["rows_contained"]
        assert st["entries_removed"] == c["entries_removed"]
        n += 1
    assert n == len(gold["chain"]) and n > 50


def test_collapse_pairs_keeps_last_row_of_a_pair():
    rows = [(0, 1, 2, 10, 8), (0, 1, 4, 10, 6), (0, 2, 1, 10, 9), (1, 0, 3, 12, 9)]
    lens = [10, 12, 20]
    kept, _ = F.filter_rows(rows, lens, contained_reads=False, max_overhang_abs=None, collapse_pairs=True)
    assert kept == [1, 2, 3]


# ---- device ------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def ov():
    from phasm_b200.overlapper import ExactOverlapper
    o = ExactOverlapper()
    yield o
    o.close()


@pytest.mark.gpu
def test_device_filter_matches_reference_chain_on_mid_cases(ov, gold):
    opts = gold["option_sets"]
    done = 0
    for key in sorted(helpers.MID_SPECS):
        cs = [c for c in gold["chain"] if c["case"] == key]
        if not cs:
            continue
        reads, l = helpers.mid_input(key)
        ov.clear()
        for name, seq in reads:
            ov.add_read(name, seq)
        rows = ov.overlaps_arrays(l)
        assert np.array_equal(rows, sort_rows(helpers.load_mid()[key]))   # canonical order already
        for c in cs:
            got, st = ov.filter_rows(**opts[c["opt"]])
            want = rows[np.asarray(expected_kept(c), dtype=np.int64)] if c["n_rows"] else rows[:0]
            assert np.array_equal(got, want.reshape(-1, 5)), (key, c["opt"])
            assert st["rows_in"] == len(rows) and st["rows_kept"] == len(want)
            assert st["rows_contained"] == c["rows_contained"]
            assert st["entries_removed"] == c["entries_removed"]
            done += 1
        # the unfiltered rows are still there
        assert np.array_equal(ov.fetch_rows(), rows)
    assert done >= 20


@pytest.mark.gpu
def test_device_filter_options_against_restatement(ov):
    from phasm_b200 import synth
    reads = synth.edge_mix(5, 200)
    ov.clear()
    lens = []
    for name, seq in reads:
        ov.add_read(name, seq)
        lens += [len(seq), len(seq)]
    rows = ov.overlaps_arrays(200)
    assert len(rows) > 0
    for opt in (dict(contained_reads=False, max_overhang_abs=None, collapse_pairs=True),
                dict(contained_reads=True, collapse_pairs=True),
                dict(contained_reads=True, min_read_length=400, min_overlap_length=250),
                dict(contained_reads=False, max_overhang_abs=0, max_overhang_rel=0.0),
                dict(contained_reads=False, max_overhang_abs=-1)):
        kept, st_w = F.filter_rows(rows.tolist(), lens, **opt)
        got, st = ov.filter_rows(**opt)
        assert np.array_equal(got, rows[np.asarray(kept, dtype=np.int64)].reshape(-1, 5)), opt
        if opt.get("contained_reads"):
            assert st["rows_contained"] == st_w["rows_contained"]
            assert st["entries_removed"] == st_w["entries_removed"]


@pytest.mark.gpu
def test_device_filter_on_empty_result(ov):
    ov.clear()
    ov.add_read("r0", "ACGTACGTACGTACGTACGT")
    assert len(ov.overlaps_arrays(1000)) == 0
    got, st = ov.filter_rows()
    assert got.shape == (0, 5) and st["rows_in"] == 0 and st["rows_kept"] == 0
