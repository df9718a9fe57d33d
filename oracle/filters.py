"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's `layout` alignment filters for
rows of the overlapper.  Only tests/ may import this.

Follows, line by line:
  phasm/cli/assembler.py:76-86,96-98   filters = [ContainedReads(), MinReadLength?, MinOverlapLength?,
                                       MaxOverhang(abs, rel)]; keep iff all(f(la) for f in filters)
  phasm/filter.py:36-57                MinReadLength        phasm/filter.py:60-75   MinOverlapLength
  phasm/filter.py:78-103               ContainedReads       phasm/filter.py:106-122 MaxOverhang
  phasm/alignments.py:237-258          get_overlap_length, get_overhang, classify

Pinned on the reference's own classes: tests/golden/filters.json (tests/golden/make_golden_filters.py).
A row is (a, b, astart, aend, bend) on entry indices with bstart == 0; `lens` = entry lengths.
"""
from typing import Iterable, List, Sequence, Tuple

A_CONTAINED, B_CONTAINED, OVERLAP_AB, OVERLAP_BA = range(4)


def classify(astart, aend, bstart, bend, la, lb) -> int:          # alignments.py:248-258
    if astart <= bstart and la - aend <= lb - bend:
        return A_CONTAINED
    if astart >= bstart and la - aend >= lb - bend:
        return B_CONTAINED
    if astart >= bstart:
        return OVERLAP_AB
    return OVERLAP_BA


def overlap_length(astart, aend, bstart, bend) -> int:            # alignments.py:237-239
    return max(aend - astart, bend - bstart)


def overhang(astart, aend, bstart, bend, la, lb) -> int:          # alignments.py:241-246
    return min(astart, bstart) + min(la - aend, lb - bend)


def filter_rows(rows: Iterable[Sequence[int]], lens: Sequence[int], contained_reads=True,
                min_read_length=0, min_overlap_length=0, max_overhang_abs=1000,
                max_overhang_rel=0.8, collapse_pairs=False) -> Tuple[List[int], dict]:
    """Indices of the rows the reference's filter chain keeps (streaming, in the given order)."""
    rows = [tuple(int(x) for x in r) for r in rows]
    removed_c, removed_l = set(), set()
    kept, n_cont = [], 0
    for i, (a, b, astart, aend, bend) in enumerate(rows):
        la, lb = lens[a], lens[b]
        ok = True
        if contained_reads:                                        # filter.py:92-103
            t = classify(astart, aend, 0, bend, la, lb)
            if t == A_CONTAINED:
                removed_c.add(a)
                ok = False
            elif t == B_CONTAINED:
                removed_c.add(b)
                ok = False
            elif a in removed_c or b in removed_c:
                ok = False
            n_cont += t in (A_CONTAINED, B_CONTAINED)
        if ok and min_read_length:                                 # filter.py:45-57 (all() short-circuits)
            if la < min_read_length:
                removed_l.add(a)
                ok = False
            elif lb < min_read_length:
                removed_l.add(b)
                ok = False
            elif a in removed_l or b in removed_l:
                ok = False
        if ok and min_overlap_length:                              # filter.py:74-75
            ok = overlap_length(astart, aend, 0, bend) >= min_overlap_length
        if ok and max_overhang_abs is not None:                    # filter.py:118-122
            thr = min(max_overhang_abs, max_overhang_rel * overlap_length(astart, aend, 0, bend))
            ok = overhang(astart, aend, 0, bend, la, lb) <= thr
        if ok:
            kept.append(i)
    if collapse_pairs:   # the row left on edge a->b when edges are added in order (assembler.py:72-74)
        last = {}
        for i in kept:
            last[(rows[i][0], rows[i][1])] = i
        kept = sorted(last.values())
    return kept, {"rows_contained": n_cont, "entries_removed": len(removed_c)}
