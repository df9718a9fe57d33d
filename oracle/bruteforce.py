"""T1 oracle: brute-force restatement of ``ExactOverlapper::overlaps``.

TEST INFRASTRUCTURE ONLY.  Nothing under ``phasm_b200/`` may import this module; only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline leg may.

It restates, as two nested loops over entry pairs, the multiset the reference's suffix-tree
DFS emits (reference: /root/reference/src/overlapper.cpp:28-150):

* R1 (overlapper.cpp:64-91) - at the leaf of every complete string B, every other string A
  that has a suffix ending on the path to that leaf reports ONLY its deepest (= longest)
  such suffix: row ``(idA, idB, |A|-L, |A|, 0, L)`` with ``L >= max(min_length, 1)``.
* R2 (overlapper.cpp:93-115) - if B as a whole (``|B| >= min_length``) occurs inside other
  strings, every occurrence ``(A, p)`` with A != B reports ``(idA, idB, p, p+|B|, 0, |B|)``.

R1 and R2 fire independently, so when B is a suffix of A the same row appears twice.
String identity is by entry index (overlapper.cpp:72-74,101-104), not by id text.
Comparison is on raw bytes (``seqan::CharString``, overlapper.h:26).

Parity pinning: this restatement is checked against the compiled reference
(``oracle/_ref``, built by ``oracle/Makefile`` from the reference's own sources) in
``tests/test_oracle.py`` and against the golden vectors under ``tests/golden/`` that
``tests/golden/make_golden.py`` generated from that compiled reference.

Pure Python, O(M^2 * len^2): only for tiny cases.
"""
from __future__ import annotations

from typing import Iterable, List, Sequence, Tuple

Row = Tuple[str, str, int, int, int, int]


def overlaps_bruteforce(entries: Sequence[Tuple[str, str]], min_length: int) -> List[Row]:
    """Return the reference's row multiset (as a sorted list) for ``entries`` = [(id, seq)]."""
    lmin = max(int(min_length), 1)
    rows: List[Row] = []
    for ia, (ida, a) in enumerate(entries):
        la = len(a)
        for ib, (idb, b) in enumerate(entries):
            if ia == ib:
                continue
            lb = len(b)
            # R1: longest suffix(a) == prefix(b) with L >= lmin (overlapper.cpp:64-91)
            for L in range(min(la, lb), lmin - 1, -1):
                if a[la - L:] == b[:L]:
                    rows.append((ida, idb, la - L, la, 0, L))
                    break
            # R2: every occurrence of whole b inside a (overlapper.cpp:93-115)
            if lb >= lmin and lb <= la:
                start = a.find(b)
                while start != -1:
                    rows.append((ida, idb, start, start + lb, 0, lb))
                    start = a.find(b, start + 1)
    rows.sort()
    return rows


_RC = bytes.maketrans(b"ACGT", b"TGCA")


def revcomp(seq: str) -> str:
    """Reverse complement of an ACGT string (what dinopy.reverse_complement does for ACGT;
    call site /root/reference/phasm/cli/assembler.py:40)."""
    return seq.encode("ascii").translate(_RC)[::-1].decode("ascii")


def both_strand_entries(reads: Iterable[Tuple[str, str]]) -> List[Tuple[str, str]]:
    """Entry list exactly as ``overlap()`` builds it (assembler.py:35-40)."""
    out: List[Tuple[str, str]] = []
    for name, seq in reads:
        out.append((name + "+", seq))
        out.append((name + "-", revcomp(seq)))
    return out
