#!/usr/bin/env python
"""Golden vectors for the f3 row filters, made with the REFERENCE's own classes
(/root/reference/phasm/filter.py, phasm/alignments.py) in this build container:

    python tests/golden/make_golden_filters.py      ->  tests/golden/filters.json

Two kinds of cases:
  * "la": random LocalAlignments with arbitrary aranges/branges (bstart != 0 too) -> the reference's
    classify(), get_overlap_length(), get_overhang(): pins the three formulas of oracle/filters.py;
  * "chain": rows of the overlapper (golden tiny / mid cases, as index rows with bstart == 0) pushed
    through the filter chain exactly as phasm/cli/assembler.py:76-86,96-98 builds and applies it,
    for several option sets -> indices of the kept rows, number of contained rows, removed entries.
"""
import json
import os
import random
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, "/root/reference")

from phasm.alignments import LocalAlignment, OrientedRead, Read, AlignmentType  # noqa: E402  (reference)
from phasm.filter import ContainedReads, MaxOverhang, MinOverlapLength, MinReadLength  # noqa: E402  (reference)

import helpers  # noqa: E402

TYPE = {AlignmentType.A_CONTAINED: 0, AlignmentType.B_CONTAINED: 1, AlignmentType.OVERLAP_AB: 2,
        AlignmentType.OVERLAP_BA: 3}

OPTION_SETS = [
    dict(contained_reads=True, min_read_length=0, min_overlap_length=0, max_overhang_abs=1000, max_overhang_rel=0.8),
    dict(contained_reads=True, min_read_length=0, min_overlap_length=0, max_overhang_abs=None, max_overhang_rel=0.8),
    dict(contained_reads=True, min_read_length=150, min_overlap_length=300, max_overhang_abs=1000, max_overhang_rel=0.8),
    dict(contained_reads=True, min_read_length=0, min_overlap_length=0, max_overhang_abs=-1, max_overhang_rel=0.8),
    dict(contained_reads=True, min_read_length=0, min_overlap_length=0, max_overhang_abs=5, max_overhang_rel=-0.5),
]


def reference_chain(rows, lens, ids, opt):
    reads = [OrientedRead(Read(i[:-1], int(l)), i[-1]) for i, l in zip(ids, lens)]
    filters = [ContainedReads()]                                  # assembler.py:76
    if opt["min_read_length"]:
        filters.append(MinReadLength(opt["min_read_length"]))     # assembler.py:78-79
    if opt["min_overlap_length"]:
        filters.append(MinOverlapLength(opt["min_overlap_length"]))
    if opt["max_overhang_abs"] is not None:
        filters.append(MaxOverhang(opt["max_overhang_abs"], opt["max_overhang_rel"]))
    kept, n_cont = [], 0
    for i, (a, b, astart, aend, bend) in enumerate(rows):
        la = LocalAlignment(reads[a], reads[b], (astart, aend), (0, bend))
        n_cont += la.classify() in (AlignmentType.A_CONTAINED, AlignmentType.B_CONTAINED)
        if all(f(la) for f in filters):                           # assembler.py:98
            kept.append(i)
    return kept, n_cont, len(filters[0].nodes_to_remove)


def main():
    rng = random.Random(20261104)
    out = {"la": [], "chain": []}
    for _ in range(400):
        la_len, lb_len = rng.randint(1, 60), rng.randint(1, 60)
        a0 = rng.randint(0, la_len - 1); a1 = rng.randint(a0 + 1, la_len)
        b0 = rng.randint(0, lb_len - 1); b1 = rng.randint(b0 + 1, lb_len)
        la = LocalAlignment(OrientedRead(Read("a", la_len), "+"), OrientedRead(Read("b", lb_len), "-"), (a0, a1), (b0, b1))
        out["la"].append([a0, a1, b0, b1, la_len, lb_len, TYPE[la.classify()], la.get_overlap_length(), la.get_overhang()])
    cases = []
    tiny = helpers.load_tiny()
    for c in tiny["tiny"][:12]:
        ids = [e[0] for e in c["entries"]]
        lens = [len(e[1]) for e in c["entries"]]
        idx = {i: k for k, i in enumerate(ids)}
        if len(idx) != len(ids):
            continue
        rows = sorted((idx[a], idx[b], s, e, be) for a, b, s, e, _, be in c["rows"])
        cases.append(("tiny", rows, lens, ids))
    mid = helpers.load_mid()
    for key in sorted(helpers.MID_SPECS):
        reads, _ = helpers.mid_input(key)
        lens, ids = [], []
        for n, s in reads:
            lens += [len(s), len(s)]
            ids += [n + "+", n + "-"]
        rows = [tuple(int(x) for x in r) for r in helpers.sort_rows(mid[key])]
        if len(rows) > 60000:
            continue
        cases.append((key, rows, lens, ids))
    for name, rows, lens, ids in cases:
        for oi, opt in enumerate(OPTION_SETS):
            kept, n_cont, n_rm = reference_chain(rows, lens, ids, opt)
            ks = set(kept)
            dropped = [i for i in range(len(rows)) if i not in ks]
            out["chain"].append({"case": name, "opt": oi, "n_rows": len(rows),
                                 # the shorter of the two index lists is stored
                                 "kept": kept if len(kept) <= len(dropped) else None,
                                 "dropped": dropped if len(dropped) < len(kept) else None,
                                 "rows_contained": n_cont, "entries_removed": n_rm,
                                 "rows": rows if name == "tiny" else None, "lens": lens if name == "tiny" else None})
    out["option_sets"] = OPTION_SETS
    with open(os.path.join(HERE, "filters.json"), "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print("la cases %d, chain cases %d" % (len(out["la"]), len(out["chain"])))


if __name__ == "__main__":
    main()
