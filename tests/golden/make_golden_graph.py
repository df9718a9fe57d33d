#!/usr/bin/env python
"""Golden vectors for the f4 assembly-graph build, made with the REFERENCE's own function
(/root/reference/phasm/assembly_graph.py:136-179 `build_assembly_graph`) in this build container:

    python tests/golden/make_golden_graph.py      ->  tests/golden/graph.json

The reference targets networkx 1.x (`g.add_edge(u, v, {attr dict})`); this image has networkx 3.x,
whose `add_edge(u, v, **attr)` rejects a positional dict.  The generator therefore gives
`AssemblyGraph.add_edge` a two-line adapter that forwards the dict as keyword attributes --
`build_assembly_graph` itself, `LocalAlignment.classify`, `get_overlap_length` and the DiGraph
"last add wins, first add keeps the adjacency slot" behaviour are the reference's / networkx's own.

Two kinds of cases:
  * "aln":   random alignments with arbitrary aranges / branges (bstart != 0, so OVERLAP_BA occurs);
  * "chain": rows of the overlapper (golden tiny / mid cases, bstart == 0) pushed through the reference's
    filter chain exactly as phasm/cli/assembler.py:76-100 does and then into build_assembly_graph.
Every case records the edges as (u, v, weight, overlap_len, rank in g.edges() order) on entry indices,
and networkx.number_of_nodes / number_of_edges.
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

import networkx  # noqa: E402

from phasm.alignments import LocalAlignment, OrientedRead, Read  # noqa: E402  (reference)
from phasm.assembly_graph import AssemblyGraph, build_assembly_graph  # noqa: E402  (reference)
from phasm.filter import ContainedReads, MaxOverhang, MinOverlapLength, MinReadLength  # noqa: E402  (reference)

import helpers  # noqa: E402


def _add_edge(self, u, v, attr_dict=None, **attr):  # networkx 1.x signature on top of networkx 3.x
    d = dict(attr_dict or {})
    d.update(attr)
    return networkx.DiGraph.add_edge(self, u, v, **d)


AssemblyGraph.add_edge = _add_edge

OPTION_SETS = [
    None,  # no filter at all
    dict(min_read_length=0, min_overlap_length=0, max_overhang_abs=1000, max_overhang_rel=0.8),
    dict(min_read_length=150, min_overlap_length=300, max_overhang_abs=1000, max_overhang_rel=0.8),
]


def graph_record(g, index):
    edges = []
    for rank, (u, v, d) in enumerate(g.edges(data=True)):
        edges.append([index[str(u)], index[str(v)], int(d["weight"]), int(d["overlap_len"]), rank])
    return {"edges": sorted(edges), "n_nodes": networkx.number_of_nodes(g), "n_edges": networkx.number_of_edges(g)}


def reference_graph(alns, lens, ids, opt):
    reads = [OrientedRead(Read(i[:-1], int(l)), i[-1]) for i, l in zip(ids, lens)]
    index = {i: k for k, i in enumerate(ids)}
    las = [LocalAlignment(reads[a], reads[b], (s, e), (bs, be)) for a, b, s, e, bs, be in alns]
    if opt is not None:
        filters = [ContainedReads()]                                  # assembler.py:76
        if opt["min_read_length"]:
            filters.append(MinReadLength(opt["min_read_length"]))
        if opt["min_overlap_length"]:
            filters.append(MinOverlapLength(opt["min_overlap_length"]))
        filters.append(MaxOverhang(opt["max_overhang_abs"], opt["max_overhang_rel"]))
        la_iter = filter(lambda x: all(f(x) for f in filters), las)   # assembler.py:98
    else:
        la_iter = iter(las)
    return graph_record(build_assembly_graph(la_iter), index)        # assembler.py:100


def main():
    rng = random.Random(20261105)
    out = {"aln": [], "chain": []}
    for case in range(40):
        n_reads = rng.randint(2, 8)
        ids = ["r%d%s" % (i, s) for i in range(n_reads) for s in "+-"]
        lens = []
        for i in range(n_reads):
            l = rng.randint(5, 60)
            lens += [l, l]
        alns = []
        for _ in range(rng.randint(1, 40)):
            a, b = rng.randrange(2 * n_reads), rng.randrange(2 * n_reads)
            if a >> 1 == b >> 1:
                continue
            a0 = rng.randint(0, lens[a] - 1); a1 = rng.randint(a0 + 1, lens[a])
            b0 = rng.randint(0, lens[b] - 1); b1 = rng.randint(b0 + 1, lens[b])
            alns.append([a, b, a0, a1, b0, b1])
        out["aln"].append({"lens": lens, "alns": alns, "graph": reference_graph(alns, lens, ids, None)})
    cases = []
    tiny = helpers.load_tiny()
    for c in tiny["tiny"][:12]:
        ids = [e[0] for e in c["entries"]]
        lens = [len(e[1]) for e in c["entries"]]
        idx = {i: k for k, i in enumerate(ids)}
        if len(idx) != len(ids) or not all(ids[2 * i][:-1] == ids[2 * i + 1][:-1] and ids[2 * i][-1] == "+"
                                           and ids[2 * i + 1][-1] == "-" for i in range(len(ids) // 2)):
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
        alns = [[a, b, s, e, 0, be] for a, b, s, e, be in rows]
        for k, opt in enumerate(OPTION_SETS):
            g = reference_graph(alns, lens, ids, opt)
            if name != "tiny":  # mid cases: a digest of the (u, v, weight, overlap_len, rank) table instead of the table
                g["sha256"] = helpers.edges_digest(g.pop("edges"))
            out["chain"].append({"case": name, "opt": opt, "n_rows": len(rows), "graph": g,
                                 "rows": rows if name == "tiny" else None, "lens": lens if name == "tiny" else None})
    with open(os.path.join(HERE, "graph.json"), "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print("aln cases %d, chain cases %d" % (len(out["aln"]), len(out["chain"])))


if __name__ == "__main__":
    main()
