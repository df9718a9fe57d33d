"""Native GFA2 reader (f4, parsing part) against golden vectors made with the reference's own
parser, and as the inverse of the GFA2 writer.  Host-only code: runs without a GPU."""
import json
import os

import numpy as np
import pytest

from phasm_b200.gfa_native import read_gfa

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "gfa_parse.json")


def parse(tmp_path, text, name="t.gfa"):
    p = tmp_path / name
    p.write_bytes(text.encode())
    return read_gfa(str(p))


def test_matches_reference_parser_on_golden_cases(tmp_path):
    with open(GOLD) as f:
        cases = json.load(f)
    assert len(cases) >= 4
    for k, c in enumerate(cases):
        names, lengths, edges = parse(tmp_path, c["text"], "c%d.gfa" % k)
        assert [[n, int(l)] for n, l in zip(names, lengths)] == c["segments"], k
        got = [[names[e["a"]] + e["a_strand"].decode(), names[e["b"]] + e["b_strand"].decode(),
                [int(e["astart"]), int(e["aend"])], [int(e["bstart"]), int(e["bend"])]] for e in edges]
        assert got == c["edges"], k


def test_python_edge_parser_matches_reference_on_golden_cases():
    from phasm.io import gfa
    with open(GOLD) as f:
        cases = json.load(f)
    n = 0
    for c in cases:
        want = iter(c["edges"])
        for line in c["text"].splitlines(keepends=True):
            if line.startswith("E"):
                sid1, sid2, ar, br, _aln, _tags = gfa.gfa2_parse_edge(line)
                assert [sid1, sid2, list(ar), list(br)] == next(want)
                n += 1
    assert n > 300
    assert gfa.gfa2_parse_edge("E\tid\ta+\tb-\t1\t2$\t3\t4\t4M\tXX:i:1\n") == ("a+", "b-", (1, 2), (3, 4), "4M", ["XX:i:1"])
    with pytest.raises(ValueError):
        gfa.gfa2_parse_edge("S\ta\t5\t*\n")


def test_errors(tmp_path):
    with pytest.raises(ValueError, match="unknown segment"):
        parse(tmp_path, "S\ta\t5\t*\nE\t*\ta+\tb-\t0\t5\t0\t5\t*\n")          # the reference: KeyError
    with pytest.raises(ValueError, match="fields"):
        parse(tmp_path, "S\ta\t5\t*\nE\t*\ta+\ta-\t0\t5\n")                    # the reference: IndexError
    with pytest.raises(ValueError, match="position"):
        parse(tmp_path, "S\ta\t5\t*\nE\t*\ta+\ta-\t0\tx\t0\t5\t*\n")           # the reference: ValueError
    with pytest.raises(ValueError, match="cannot open"):
        read_gfa(str(tmp_path / "missing.gfa"))


def test_inverse_of_the_python_writer(tmp_path):
    from phasm.io import gfa
    rng = np.random.default_rng(3)
    names = ["read/%d" % i for i in range(25)]
    lens = rng.integers(100, 500, size=25)
    rows = []
    lines = [gfa.gfa_header()] + [gfa.gfa_line("S", n, int(l), "*") for n, l in zip(names, lens)]
    for _ in range(200):
        a, b = rng.choice(25, size=2, replace=False)
        L = int(rng.integers(10, min(lens[a], lens[b])))
        sa, sb = "+-"[int(rng.integers(2))], "+-"[int(rng.integers(2))]
        rows.append((int(a), sa, int(b), sb, int(lens[a]) - L, int(lens[a]), 0, L))
        lines.append(gfa.gfa_line("E", "*", names[a] + sa, names[b] + sb, int(lens[a]) - L, int(lens[a]), 0, L, "*"))
    got_names, got_lens, edges = parse(tmp_path, "".join(lines))
    assert got_names == names and got_lens.tolist() == lens.tolist()
    got = [(int(e["a"]), e["a_strand"].decode(), int(e["b"]), e["b_strand"].decode(), int(e["astart"]),
            int(e["aend"]), int(e["bstart"]), int(e["bend"])) for e in edges]
    assert got == rows


@pytest.mark.gpu
def test_round_trip_of_the_native_writer_on_device_rows(tmp_path):
    """FASTA -> overlaps on the GPU -> ovl_write_gfa -> ovl_gfa_read gives back the rows."""
    from phasm_b200 import synth
    from phasm_b200.overlapper import ExactOverlapper
    sh = synth.CONFIGS[1].scaled(0.01)
    buf, off, names = synth.generate(sh)
    ov = ExactOverlapper()
    try:
        ov.add_reads_bulk(buf, off, names)
        rows = ov.overlaps_arrays(sh.min_length)
        assert len(rows) > 100
        path = tmp_path / "out.gfa"
        with open(path, "wb") as f:
            ov.write_gfa(f)
    finally:
        ov.close()
    seg_names, seg_lens, edges = read_gfa(str(path))
    assert seg_names == list(names) and set(seg_lens.tolist()) == {sh.read_len}
    # entry index = 2 * read + strand
    a = 2 * edges["a"].astype(np.int64) + (edges["a_strand"] == b"-")
    b = 2 * edges["b"].astype(np.int64) + (edges["b_strand"] == b"-")
    got = np.stack([a, b, edges["astart"], edges["aend"], edges["bend"]], axis=1).astype(np.uint32)
    assert np.array_equal(got, rows) and not edges["bstart"].any()
