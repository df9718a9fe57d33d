#!/usr/bin/env python
"""Golden vectors for the native GFA2 reader (f4), made with the REFERENCE's own parser
(/root/reference/phasm/io/gfa.py: gfa2_parse_segments, gfa2_parse_edge) in this build container:

    python tests/golden/make_golden_gfa.py      ->  tests/golden/gfa_parse.json

Each case is a GFA2 text and what the reference's two passes return for it: the segment dict
(name -> length, in dict order) and the parsed edges (sid1, sid2, arange, brange).
"""
import io
import json
import os
import random
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, "/root/reference")

from phasm.io import gfa  # noqa: E402  (the reference)


def reference_parse(text):
    reads = gfa.gfa2_parse_segments(io.StringIO(text))                       # assembler.py:58-59
    segs = [[r.id, len(r)] for r in reads.values()]
    edges = []
    for line in io.StringIO(text):                                          # assembler.py:93-95
        if line.startswith('E'):
            sid1, sid2, ar, br, _aln, _tags = gfa.gfa2_parse_edge(line)
            _ = reads[sid1[:-1]], reads[sid2[:-1]]                          # gfa2_line_to_la lookups
            edges.append([sid1, sid2, list(ar), list(br)])
    return segs, edges


def main():
    rng = random.Random(4)
    cases = []
    # what `phasm overlap` writes (assembler.py:30,38,46-48)
    lines = [gfa.gfa_header()]
    names = ["r%03d" % i for i in range(40)]
    lens = [rng.randint(50, 400) for _ in names]
    for n, l in zip(names, lens):
        lines.append(gfa.gfa_line("S", n, l, "*"))
    for _ in range(300):
        a, b = rng.sample(range(40), 2)
        L = rng.randint(10, min(lens[a], lens[b]))
        lines.append(gfa.gfa_line("E", "*", names[a] + rng.choice("+-"), names[b] + rng.choice("+-"),
                                  lens[a] - L, lens[a], 0, L, "*"))
    cases.append("".join(lines))
    # oddities the reference parser accepts
    cases.append("H\tVN:z:2.0\n"
                 "E\t*\tx+\ty-\t5\t10$\t0\t5\t*\n"            # edge before its segments
                 "S\tx\t10\t*\n"
                 "S\ty\t7\tACGTACG\n"
                 "S\tx\t12\t*\n"                               # a repeated name: the later length wins
                 "E\tid7\ty+\tx+\t0$\t7$\t 3\t10 \t4M3D\tXX:i:1\tYY:Z:a b\n"
                 "E\t*\t x- \t y+ \t1\t2\t3\t4\t*   \r\n"
                 "# comment\n"
                 "F\tx\tfrag+\t0\t5\t0\t5\t*\n"
                 "S\tz z\t+5\t*\n"
                 "E\t*\tz z+\tx-\t0\t5$\t7\t12$\t*")           # no newline at the end
    cases.append("")
    cases.append("S\tonly\t3\t*\n")
    out = []
    for text in cases:
        segs, edges = reference_parse(text)
        out.append({"text": text, "segments": segs, "edges": edges})
    with open(os.path.join(HERE, "gfa_parse.json"), "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print("cases", len(out), "edges", [len(c["edges"]) for c in out])


if __name__ == "__main__":
    main()
