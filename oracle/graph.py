"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's build_assembly_graph for rows of the
overlapper (graph part).  Only tests/ may import this.

Follows phasm/assembly_graph.py:136-179 (two add_edge calls per OVERLAP_AB / OVERLAP_BA alignment,
nothing for the contained types), phasm/alignments.py:237-258 (classify, get_overlap_length) and the
networkx.DiGraph behaviour the reference relies on: adding an existing edge replaces its attributes
and keeps its place in the adjacency order; a node exists once it ends an edge.

Pinned on the reference's own function: tests/golden/graph.json (tests/golden/make_golden_graph.py).
An alignment is (a, b, astart, aend, bstart, bend) on entry indices; entry e ^ 1 is the other strand
of entry e (OrientedRead.reverse()); `lens` = entry lengths.
"""
from typing import Dict, Iterable, List, Sequence, Tuple

from .filters import OVERLAP_AB, OVERLAP_BA, classify, overlap_length


def build_graph(alns: Iterable[Sequence[int]], lens: Sequence[int]) -> Tuple[List[Tuple[int, int, int, int, int]], int]:
    """-> (edges sorted by (u, v) as (u, v, edge_len, overlap_len, first_seq), number of nodes).

    first_seq = 2 * (index of the alignment that first added the edge) + (0 | 1: which of its calls)."""
    edges: Dict[Tuple[int, int], List[int]] = {}
    nodes = set()

    def add_edge(u, v, w, ol, seq):
        nodes.add(u)
        nodes.add(v)
        e = edges.get((u, v))
        if e is None:
            edges[(u, v)] = [w, ol, seq]
        else:
            e[0], e[1] = w, ol

    for i, (a, b, astart, aend, bstart, bend) in enumerate(alns):
        la, lb = int(lens[a]), int(lens[b])
        t = classify(astart, aend, bstart, bend, la, lb)
        ol = overlap_length(astart, aend, bstart, bend)
        if t == OVERLAP_AB:                                        # assembly_graph.py:147-156
            add_edge(a, b, astart - bstart, ol, 2 * i)
            add_edge(b ^ 1, a ^ 1, (lb - bend) - (la - aend), ol, 2 * i + 1)
        elif t == OVERLAP_BA:                                      # assembly_graph.py:163-172
            add_edge(b, a, bstart - astart, ol, 2 * i)
            add_edge(a ^ 1, b ^ 1, (la - aend) - (lb - bend), ol, 2 * i + 1)
    out = sorted((u, v, w, ol, seq) for (u, v), (w, ol, seq) in edges.items())
    return out, len(nodes)
