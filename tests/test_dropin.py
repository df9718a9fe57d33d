"""Drop-in surface: the ``phasm`` package (extension module, GFA2 writer, FASTA reader, CLI).

CPU tests check the host logic and the module surface; the ``gpu`` tests run the CLI end to
end on the toy FASTA of the reference's own fixtures (tests/data/test.fasta: blank lines, one
wrapped record) against the golden rows the compiled reference produced for it.
"""
import io
import os
import subprocess
import sys

import pytest

import helpers

ROOT = helpers.ROOT

TOY_FASTA = (">read1\nATGCATG\n\n>read2\nCATGGGGGGGG\n\n>read3\nGGGGGGG\nACTGCTAC\n\n"
             ">read4\nGCTACGTGTGTGTG\n\n")


@pytest.fixture(scope="module")
def ext():
    import __graft_entry__ as g
    g.build_library()
    g.build_pybind()
    import phasm.overlapper as mod
    return mod


def write_toy(tmp_path):
    p = tmp_path / "toy.fasta"
    p.write_text(TOY_FASTA)
    return str(p)


def expected_gfa(toy):
    from phasm.io import gfa
    head = gfa.gfa_header() + "".join(gfa.gfa_line("S", n, len(s), "*") for n, s in toy["reads"])
    edges = sorted(gfa.gfa_line("E", "*", *r, "*") for r in toy["rows"])
    return head, edges


def split_gfa(text):
    lines = text.splitlines(keepends=True)
    head = "".join(l for l in lines if not l.startswith("E"))
    return head, sorted(l for l in lines if l.startswith("E"))


# ---- CPU ------------------------------------------------------------------------------------
def test_gfa_lines_are_byte_exact():
    from phasm.io import gfa
    assert gfa.gfa_header() == "H\tVN:z:2.0\n"
    assert gfa.gfa_header("1.0") == "H\tVN:z:1.0\n"
    assert gfa.gfa_header(trace_spacing=100) == "H\tVN:z:2.0\tTS:i:100\n"
    assert gfa.gfa_line("S", "r1", 7, "*") == "S\tr1\t7\t*\n"
    assert gfa.gfa_line("E", "*", "a+", "b-", 3, 7, 0, 4, "*") == "E\t*\ta+\tb-\t3\t7\t0\t4\t*\n"
    assert gfa.gfa_line() == "\n"


def test_fasta_reader_matches_toy_fixture(tmp_path):
    from phasm.io.fasta import FastaReader, reverse_complement
    toy = helpers.load_tiny()["toy"]
    got = [(e.name.decode(), e.sequence.decode(), e.length)
           for e in FastaReader(write_toy(tmp_path)).entries()]
    assert got == [(n, s, len(s)) for n, s in toy["reads"]]
    assert reverse_complement("ATGCATG") == helpers.revcomp("ATGCATG")
    assert reverse_complement(b"AACG") == b"CGTT"
    # no trailing newline, CRLF, header with spaces
    p = tmp_path / "x.fa"
    p.write_bytes(b">a b  \r\nAC\r\nGT\r\n>c\nTT")
    got = [(e.name, e.sequence) for e in FastaReader(str(p)).entries()]
    assert got == [(b"a b", b"ACGT"), (b"c", b"TT")]


def test_extension_module_surface(ext):
    assert ext.__doc__ == "Optimised C++ all-vs-all exact overlap finder for PHASM"
    cls = ext.ExactOverlapper
    # positional-only names arg0/arg1 like the reference binding (src/phasm.cpp:14-15)
    assert "arg0: str, arg1: str" in cls.add_sequence.__doc__
    assert "overlaps(self: phasm.overlapper.ExactOverlapper, arg0: " in cls.overlaps.__doc__
    for extra in ("add_read", "add_fasta", "overlaps_arrays", "write_gfa", "stats", "filter_rows"):
        assert hasattr(cls, extra)


@pytest.mark.skipif(helpers.has_gpu(), reason="a GPU is present")
def test_extension_has_no_cpu_fallback(ext):
    with pytest.raises(RuntimeError) as ei:
        ext.ExactOverlapper()
    assert "no CPU fallback" in str(ei.value)


def test_cli_parser_matches_reference_flags(tmp_path, capsys):
    from phasm.cli import assembler
    with pytest.raises(SystemExit):
        assembler.main(["overlap"])  # fasta_input is required
    assert assembler.main([]) == 0  # no subcommand: prints help like the reference
    assert "overlap" in capsys.readouterr().out
    assert assembler.main(["layout", "x.gfa"]) == 2  # other stages are not part of this build


# ---- GPU ------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_extension_reference_calls_on_toy(ext):
    toy = helpers.load_tiny()["toy"]
    ov = ext.ExactOverlapper()
    for name, seq in toy["reads"]:
        ov.add_sequence(name + "+", seq)
        ov.add_sequence(name + "-", helpers.revcomp(seq).encode())  # bytes accepted too
    rows = ov.overlaps(toy["min_length"])
    assert isinstance(rows, list) and all(isinstance(r, tuple) and len(r) == 6 for r in rows)
    assert sorted(rows) == sorted(toy["rows"])
    assert ov.overlaps(toy["min_length"]) == rows  # repeatable, fresh list
    with pytest.raises(TypeError):
        ov.overlaps(-1)
    with pytest.raises(TypeError):
        ov.overlaps(3.5)
    ov.add_sequence("n", "ACGN")
    with pytest.raises(ValueError):
        ov.overlaps(3)


@pytest.mark.gpu
def test_extension_empty_set_returns_no_rows(ext):
    assert ext.ExactOverlapper().overlaps(1000) == []


@pytest.mark.gpu
@pytest.mark.parametrize("route", ["native", "python"])
def test_cli_overlap_on_toy_fasta(ext, tmp_path, route):
    toy = helpers.load_tiny()["toy"]
    fa = write_toy(tmp_path)
    out = tmp_path / ("out_%s.gfa" % route)
    cmd = [sys.executable, "-m", "phasm", "overlap", "-l", str(toy["min_length"]), "-o", str(out)]
    if route == "python":
        cmd.append("--python-io")
    r = subprocess.run(cmd + [fa], cwd=ROOT, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    for msg in ("Building suffix tree and searching for pairwise overlaps...", "Writing to GFA2...", "Done."):
        assert msg in r.stderr
    head, edges = split_gfa(out.read_text())
    want_head, want_edges = expected_gfa(toy)
    assert head == want_head           # header + S-lines byte-exact, in FASTA order
    assert edges == want_edges         # E-lines as a sorted multiset


@pytest.mark.gpu
def test_cli_routes_agree_on_synthetic_reads(ext, tmp_path):
    from phasm_b200 import synth
    sh = synth.CONFIGS[1].scaled(0.004)
    buf, off, names = synth.generate(sh)
    fa = tmp_path / "syn.fasta"
    with open(fa, "w") as f:
        for i, n in enumerate(names):
            s = buf[int(off[i]):int(off[i + 1])].tobytes().decode()
            # wrap every other record over several lines
            if i % 2:
                s = "\n".join(s[j:j + 997] for j in range(0, len(s), 997))
            f.write(">%s\n%s\n" % (n, s))
    outs = []
    for extra in ([], ["--python-io"]):
        out = tmp_path / ("o%d.gfa" % len(outs))
        r = subprocess.run([sys.executable, "-m", "phasm", "overlap", "-l", str(sh.min_length),
                            "-o", str(out)] + extra + [str(fa)], cwd=ROOT, capture_output=True,
                           text=True, timeout=600)
        assert r.returncode == 0, r.stderr
        outs.append(out.read_bytes())
    assert outs[0] == outs[1]          # both routes sort by entry index: identical bytes
    assert outs[0].count(b"\nE\t") > 100


@pytest.mark.gpu
def test_cli_stdout_default(ext, tmp_path):
    toy = helpers.load_tiny()["toy"]
    r = subprocess.run([sys.executable, "-m", "phasm", "overlap", "-l", "3", write_toy(tmp_path)],
                       cwd=ROOT, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    head, edges = split_gfa(r.stdout)
    want_head, want_edges = expected_gfa(toy)
    assert head == want_head and edges == want_edges


@pytest.mark.gpu
def test_extension_filter_rows_matches_restatement(ext):
    """The additive filter_rows of the drop-in module (ovl_filter_rows) against the CPU restatement
    of the reference's layout filters (oracle/filters.py, pinned on the reference's own classes)."""
    import numpy as np
    from oracle import filters as F
    from phasm_b200 import synth
    reads = synth.edge_mix(9, 200)
    ov = ext.ExactOverlapper()
    lens = []
    for name, seq in reads:
        ov.add_read(name, seq)
        lens += [len(seq), len(seq)]
    rows = ov.overlaps_arrays(200)
    assert len(rows) > 0
    for opt in (dict(), dict(contained_reads=True, collapse_pairs=True, max_overhang_abs=None),
                dict(contained_reads=False, min_overlap_length=400)):
        kept, _ = F.filter_rows(rows.tolist(), lens, **opt)
        got = ov.filter_rows(**opt)
        assert np.array_equal(got, rows[np.asarray(kept, dtype=np.int64)].reshape(-1, 5)), opt
