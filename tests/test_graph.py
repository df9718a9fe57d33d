"""f4 assembly-graph build: the CPU restatement against golden vectors made with the
reference's own build_assembly_graph, and the device post-pass (ovl_build_graph) against both."""
import json
import os

import numpy as np
import pytest

import helpers
from helpers import edges_digest, first_seq_to_rank, sort_rows
from oracle import filters as F
from oracle import graph as G

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "graph.json")


@pytest.fixture(scope="module")
def gold():
    with open(GOLD) as f:
        return json.load(f)


def filter_kwargs(opt):
    return dict(contained_reads=True, **opt)


def chain_input(c):
    if c["case"] == "tiny":
        return [tuple(r) for r in c["rows"]], c["lens"], None
    reads, l = helpers.mid_input(c["case"])
    lens = []
    for _, s in reads:
        lens += [len(s), len(s)]
    rows = [tuple(int(x) for x in r) for r in sort_rows(helpers.load_mid()[c["case"]])]
    return rows, lens, (reads, l)


def check_graph(edges, n_nodes, want, what):
    """edges: (u, v, w, ol, first_seq) sorted by (u, v)."""
    assert n_nodes == want["n_nodes"], what
    assert len(edges) == want["n_edges"], what
    ranked = first_seq_to_rank(edges)
    if "edges" in want:
        assert [list(e) for e in ranked] == want["edges"], what
    else:
        assert edges_digest(ranked) == want["sha256"], what


def test_restatement_matches_reference_on_general_alignments(gold):
    n_ba = 0
    for c in gold["aln"]:
        edges, n_nodes = G.build_graph(c["alns"], c["lens"])
        check_graph(edges, n_nodes, c["graph"], "aln")
        n_ba += sum(F.classify(s, e, bs, be, c["lens"][a], c["lens"][b]) == F.OVERLAP_BA for a, b, s, e, bs, be in c["alns"])
    assert n_ba > 20  # the OVERLAP_BA branch (assembly_graph.py:163-172) is exercised


def test_restatement_matches_reference_layout_chain(gold):
    n = 0
    for c in gold["chain"]:
        rows, lens, _ = chain_input(c)
        if c["opt"] is not None:
            kept, _ = F.filter_rows(rows, lens, **filter_kwargs(c["opt"]))
            rows = [rows[i] for i in kept]
        edges, n_nodes = G.build_graph([(a, b, s, e, 0, be) for a, b, s, e, be in rows], lens)
        check_graph(edges, n_nodes, c["graph"], (c["case"], c["opt"]))
        n += 1
    assert n == len(gold["chain"]) and n >= 30


def test_last_alignment_of_a_pair_decides_the_attributes():
    lens = [100, 100, 100, 100]
    alns = [(0, 2, 40, 100, 0, 60), (0, 2, 50, 100, 0, 50)]
    edges, n_nodes = G.build_graph(alns, lens)
    assert edges == [(0, 2, 50, 50, 0), (3, 1, 50, 50, 1)] and n_nodes == 4


@pytest.mark.skipif(not os.path.isdir("/root/reference/phasm"), reason="reference checkout not mounted")
def test_restatement_against_the_live_reference_function():
    """oracle/graph.py against build_assembly_graph itself (imported from /root/reference in a
    subprocess, because its package is also called `phasm`), on fresh random alignments."""
    import subprocess
    import sys
    import tempfile
    rng = np.random.default_rng(20261105)
    cases = []
    for _ in range(60):
        n_reads = int(rng.integers(2, 7))
        lens = []
        for _ in range(n_reads):
            l = int(rng.integers(4, 40))
            lens += [l, l]
        alns = []
        for _ in range(int(rng.integers(1, 30))):
            a, b = (int(x) for x in rng.integers(0, 2 * n_reads, size=2))
            if a >> 1 == b >> 1:
                continue
            a0 = int(rng.integers(0, lens[a])); a1 = int(rng.integers(a0 + 1, lens[a] + 1))
            b0 = int(rng.integers(0, lens[b])); b1 = int(rng.integers(b0 + 1, lens[b] + 1))
            alns.append([a, b, a0, a1, b0, b1])
        cases.append({"lens": lens, "alns": alns})
    script = r'''
import json, sys
sys.path.insert(0, "/root/reference")
import networkx
from phasm.alignments import LocalAlignment, OrientedRead, Read
from phasm.assembly_graph import AssemblyGraph, build_assembly_graph
def _add_edge(self, u, v, attr_dict=None, **attr):      # networkx 1.x call on networkx >= 2
    d = dict(attr_dict or {}); d.update(attr)
    return networkx.DiGraph.add_edge(self, u, v, **d)
AssemblyGraph.add_edge = _add_edge
out = []
for c in json.load(open(sys.argv[1])):
    ids = ["r%d%s" % (i // 2, "+-"[i % 2]) for i in range(len(c["lens"]))]
    reads = [OrientedRead(Read(i[:-1], int(l)), i[-1]) for i, l in zip(ids, c["lens"])]
    index = {i: k for k, i in enumerate(ids)}
    g = build_assembly_graph(LocalAlignment(reads[a], reads[b], (s, e), (bs, be)) for a, b, s, e, bs, be in c["alns"])
    out.append({"n_nodes": g.number_of_nodes(),
                "edges": sorted([index[str(u)], index[str(v)], int(d["weight"]), int(d["overlap_len"]), r]
                                for r, (u, v, d) in enumerate(g.edges(data=True)))})
json.dump(out, open(sys.argv[2], "w"))
'''
    with tempfile.TemporaryDirectory() as td:
        fin, fout, fs = os.path.join(td, "in.json"), os.path.join(td, "out.json"), os.path.join(td, "run.py")
        with open(fin, "w") as f:
            json.dump(cases, f)
        with open(fs, "w") as f:
            f.write(script)
        r = subprocess.run([sys.executable, "-W", "ignore", fs, fin, fout], capture_output=True, text=True, timeout=300, cwd=td)
        if r.returncode != 0 and "No module named" in r.stderr:
            pytest.skip("reference dependencies missing: " + r.stderr.strip().splitlines()[-1])
        assert r.returncode == 0, r.stderr[-2000:]
        with open(fout) as f:
            want = json.load(f)
    for c, w in zip(cases, want):
        edges, n_nodes = G.build_graph(c["alns"], c["lens"])
        assert n_nodes == w["n_nodes"]
        assert [list(e) for e in first_seq_to_rank(edges)] == w["edges"]


def test_host_graph_object_reproduces_networkx_insertion_order(gold):
    """phasm.assembly_graph.edges_to_digraph on the restatement's edge list: same nodes, edges,
    attributes and g.edges() order as the reference's graph (the golden `rank` column)."""
    networkx = pytest.importorskip("networkx")
    from phasm.assembly_graph import edges_to_digraph
    from phasm_b200.overlapper import EDGE_DTYPE
    n = 0
    for c in gold["aln"]:
        edges, n_nodes = G.build_graph(c["alns"], c["lens"])
        arr = np.array(edges, dtype=EDGE_DTYPE) if edges else np.zeros(0, dtype=EDGE_DTYPE)
        ids = ["r%d%s" % (i // 2, "+-"[i % 2]) for i in range(len(c["lens"]))]
        g = edges_to_digraph(arr, ids)
        assert networkx.number_of_nodes(g) == c["graph"]["n_nodes"] and networkx.number_of_edges(g) == c["graph"]["n_edges"]
        index = {name: i for i, name in enumerate(ids)}
        got = sorted([index[u], index[v], d["weight"], d["overlap_len"], rank]
                     for rank, (u, v, d) in enumerate(g.edges(data=True)))
        assert got == c["graph"]["edges"]
        n += 1
    assert n == len(gold["aln"])


# ---- device ------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def ov():
    from phasm_b200.overlapper import ExactOverlapper
    o = ExactOverlapper()
    yield o
    o.close()


def device_edges(arr):
    return [(int(e["u"]), int(e["v"]), int(e["edge_len"]), int(e["overlap_len"]), int(e["first_seq"])) for e in arr]


@pytest.mark.gpu
def test_device_graph_matches_reference_on_mid_cases(ov, gold):
    done = 0
    for key in sorted(helpers.MID_SPECS):
        cs = [c for c in gold["chain"] if c["case"] == key]
        if not cs:
            continue
        reads, l = helpers.mid_input(key)
        ov.clear()
        lens = []
        for name, seq in reads:
            ov.add_read(name, seq)
            lens += [len(seq), len(seq)]
        rows = ov.overlaps_arrays(l)
        for c in cs:
            if c["opt"] is None:
                arr, st = ov.build_graph(filtered=False)
                n_in = len(rows)
            else:
                kept, _ = ov.filter_rows(**filter_kwargs(c["opt"]))
                arr, st = ov.build_graph(filtered=True)
                n_in = len(kept)
            edges = device_edges(arr)
            assert edges == sorted(edges)
            check_graph(edges, st["n_nodes"], c["graph"], (key, c["opt"]))
            assert st["rows_in"] == n_in and st["n_edges"] == len(edges)
            done += 1
    assert done >= 15


@pytest.mark.gpu
def test_device_graph_against_restatement_on_edge_mix(ov):
    from phasm_b200 import synth
    reads = synth.edge_mix(7, 150)
    ov.clear()
    lens = []
    for name, seq in reads:
        ov.add_read(name, seq)
        lens += [len(seq), len(seq)]
    rows = ov.overlaps_arrays(150)
    assert len(rows) > 0
    for opt in (None, dict(contained_reads=True), dict(contained_reads=False, max_overhang_abs=None),
                dict(contained_reads=True, collapse_pairs=True)):
        if opt is None:
            use = rows.tolist()
            arr, st = ov.build_graph(filtered=False)
        else:
            kept, _ = F.filter_rows(rows.tolist(), lens, **opt)
            use = [rows[i].tolist() for i in kept]
            ov.filter_rows(**opt)
            arr, st = ov.build_graph(filtered=True)
        want, n_nodes = G.build_graph([(a, b, s, e, 0, be) for a, b, s, e, be in use], lens)
        assert device_edges(arr) == want, opt
        assert st["n_nodes"] == n_nodes and st["rows_in"] == len(use)


@pytest.mark.gpu
def test_device_graph_needs_a_filter_result_and_pairs(ov):
    ov.clear()
    ov.add_read("r0", "ACGTACGTACGTACGTACGTAAAA")
    ov.add_read("r1", "ACGTACGTACGTAAAACCCCGGGG")
    ov.overlaps_arrays(8)
    with pytest.raises(ValueError):
        ov.build_graph(filtered=True)          # no filter_rows call for this pass
    arr, st = ov.build_graph(filtered=False)
    assert st["rows_in"] > 0
    ov.clear()
    ov.add_sequence("x", "ACGTACGTACGTACGTACGTAAAA")   # not strand-paired
    ov.add_sequence("y", "ACGTACGTACGTAAAACCCCGGGG")
    if len(ov.overlaps_arrays(8)):
        with pytest.raises(ValueError):
            ov.build_graph(filtered=False)


@pytest.mark.gpu
def test_device_graph_of_a_pass_without_rows_is_empty(ov):
    """No overlaps -> the reference's build_assembly_graph returns an empty graph
    (phasm/assembly_graph.py:136-179 adds nothing); filter_rows + build_graph(filtered=True) must
    not raise on that input."""
    ov.clear()
    ov.add_read("r0", "ACGTACGTACGTACGTACGT")
    assert len(ov.overlaps_arrays(1000)) == 0
    got, fst = ov.filter_rows()
    assert got.shape == (0, 5) and fst["rows_kept"] == 0
    edges, st = ov.build_graph(filtered=True)
    assert len(edges) == 0 and st["rows_in"] == 0 and st["n_edges"] == 0 and st["n_nodes"] == 0


@pytest.mark.gpu
def test_drop_in_module_drives_overlap_filter_and_graph(gold):
    """f3 -> f4 through the drop-in extension module itself (`phasm.overlapper.ExactOverlapper`, the object
    the reference's `overlap` stage creates: src/phasm.cpp:8-18): overlaps, the layout filters, then
    `phasm.assembly_graph.build_assembly_graph` (reference: phasm/assembly_graph.py:136-179) -- against the
    golden graphs made with the reference's own function (tests/golden/graph.json)."""
    import __graft_entry__ as g
    g.build_library()
    g.build_pybind()
    import phasm.overlapper as mod
    from phasm import assembly_graph as AG
    done = 0
    for key in sorted(helpers.MID_SPECS)[:3]:
        cs = [c for c in gold["chain"] if c["case"] == key and c["opt"] is not None]
        if not cs:
            continue
        reads, l = helpers.mid_input(key)
        ext = mod.ExactOverlapper()
        for name, seq in reads:
            ext.add_read(name, seq)
        rows = ext.overlaps_arrays(l)
        assert len(rows) == len(helpers.load_mid()[key])
        assert len(ext) == 2 * len(reads) and ext.ids()[:2] == [reads[0][0] + "+", reads[0][0] + "-"]
        assert ext.lengths().tolist() == [len(s) for _, s in reads for _ in (0, 1)]
        for c in cs:
            kept = ext.filter_rows(**filter_kwargs(c["opt"]))
            arr, st = ext.build_graph(filtered=True)
            edges = device_edges(arr)
            assert edges == sorted(edges) and st["rows_in"] == len(kept)
            check_graph(edges, st["n_nodes"], c["graph"], (key, c["opt"]))
            dg = AG.build_assembly_graph(ext, filtered=True)
            assert dg.number_of_nodes() == c["graph"]["n_nodes"] and dg.number_of_edges() == c["graph"]["n_edges"]
            ids = ext.ids()
            first = min(edges, key=lambda e: e[4]) if edges else None
            if first:
                u, v = next(iter(dg.edges()))
                assert (u, v) == (ids[first[0]], ids[first[1]])      # networkx insertion order
                assert dg[u][v] == {"weight": first[2], "overlap_len": first[3]}
            done += 1
    assert done >= 3
