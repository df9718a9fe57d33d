"""``build_assembly_graph`` for the drop-in package (graph part).

The reference builds its ``AssemblyGraph`` (a ``networkx.DiGraph``) from a stream of
``LocalAlignment`` objects, one Python ``add_edge`` pair per overlap
(/root/reference/phasm/assembly_graph.py:136-179, fed by phasm/cli/assembler.py:93-100).  Here the
edge list comes from the device (``ExactOverlapper.build_graph`` = ``ovl_build_graph``: classify,
the two directed edges per OVERLAP row, "last alignment of a node pair wins") and this module only
turns it into the graph object a consumer of the reference expects: nodes are the oriented-read
names (``"<read id><strand>"``, what ``str(OrientedRead)`` gives), edge attributes ``weight``
(edge_len) and ``overlap_len``, nodes and adjacency lists in the reference's insertion order.

networkx is needed only by ``edges_to_digraph``; ``edges_in_insertion_order`` is plain numpy.
"""
from __future__ import annotations

from typing import Iterable, List, Sequence, Tuple

import numpy as np


def edges_in_insertion_order(edges: np.ndarray) -> np.ndarray:
    """The device edge array (sorted by (u, v)) in the order the reference's ``add_edge`` calls first
    created the edges (``first_seq``): adding them in this order reproduces networkx's node order and
    adjacency order of the reference's graph."""
    return edges[np.argsort(edges["first_seq"], kind="stable")]


def edge_tuples(edges: np.ndarray, ids: Sequence[str], edge_len: str = "weight",
                overlap_len: str = "overlap_len") -> List[Tuple[str, str, dict]]:
    """``[(u_name, v_name, {edge_len: .., overlap_len: ..})]`` in insertion order."""
    out = []
    for e in edges_in_insertion_order(edges):
        out.append((ids[int(e["u"])], ids[int(e["v"])],
                    {edge_len: int(e["edge_len"]), overlap_len: int(e["overlap_len"])}))
    return out


def edges_to_digraph(edges: np.ndarray, ids: Sequence[str], edge_len: str = "weight",
                     overlap_len: str = "overlap_len"):
    """A ``networkx.DiGraph`` equal to the reference's ``build_assembly_graph`` result on the same
    alignments (nodes named like ``str(OrientedRead)``; graph attributes ``edge_len`` /
    ``overlap_len`` as ``AssemblyGraph.__init__`` stores them)."""
    import networkx
    g = networkx.DiGraph(edge_len=edge_len, overlap_len=overlap_len)
    for u, v, d in edge_tuples(edges, ids, edge_len, overlap_len):
        g.add_edge(u, v, **d)
    return g


def build_assembly_graph(overlapper, filtered: bool = True, edge_len: str = "weight",
                         overlap_len: str = "overlap_len"):
    """The assembly graph of the rows of ``overlapper``'s last ``overlaps*`` call (the rows kept by its
    last ``filter_rows`` call when ``filtered``): ``phasm_b200.overlapper.ExactOverlapper.build_graph``
    on the device, then ``edges_to_digraph``."""
    edges, _ = overlapper.build_graph(filtered=filtered)
    return edges_to_digraph(edges, overlapper.ids(), edge_len, overlap_len)
