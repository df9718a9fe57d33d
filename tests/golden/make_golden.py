#!/usr/bin/env python
"""Generate the golden vectors under tests/golden/ from the COMPILED REFERENCE (oracle/_ref).

Run in the build container (where /root/reference is mounted and `make -C oracle ref` works):

    python tests/golden/make_golden.py

Every vector stores the inputs (or the deterministic generator parameters) and the rows the
unmodified reference overlapper returned for them, driven exactly as
/root/reference/phasm/cli/assembler.py:39-42 drives it.  The reference publishes no golden
vectors for this path, so these files are what pins the oracles
(oracle/bruteforce.py, oracle/overlap_oracle.c) and, through them, the CUDA path.
"""
from __future__ import annotations

import json
import os
import random
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref  # noqa: E402
from oracle.bruteforce import both_strand_entries, revcomp  # noqa: E402
from phasm_b200 import synth  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))

# the four toy reads of the reference's only FASTA fixture (tests/data/test.fasta)
TOY_READS = [("read1", "ATGCATG"), ("read2", "CATGGGGGGGG"),
             ("read3", "GGGGGGGACTGCTAC"), ("read4", "GCTACGTGTGTGTG")]


def tiny_cases(seed: int, n_cases: int):
    rnd = random.Random(seed)
    cases = []
    while len(cases) < n_cases:
        alpha = rnd.choice(["A", "AC", "ACGT", "ACGT"])
        ents = []
        for i in range(rnd.randint(1, 7)):
            s = "".join(rnd.choice(alpha) for _ in range(rnd.randint(0, 14)))
            if ents and rnd.random() < 0.35:
                t = rnd.choice(ents)[1]
                if t:
                    i0 = rnd.randrange(len(t))
                    i1 = rnd.randint(i0, len(t))
                    s = rnd.choice([t[i0:i1] + s, s + t[i0:i1], t[i0:i1], t])
            ents.append(("e%d" % i, s))
        if all(not s for _, s in ents):
            continue  # the reference segfaults when every entry is empty
        if rnd.random() < 0.5:
            ents = both_strand_entries(ents)
        l = rnd.randint(0, 6)
        rows = ref.ref_overlaps(ents, l)
        cases.append({"entries": ents, "min_length": l, "rows": rows})
    return cases


def index_rows(rows, ids):
    """tuples -> (n,5) uint32 on entry indices (ids are unique in the mid-size sets)."""
    pos = {s: i for i, s in enumerate(ids)}
    arr = np.array([(pos[a], pos[b], s, e, be) for a, b, s, e, _, be in rows], dtype=np.uint32)
    if len(arr) == 0:
        arr = arr.reshape(0, 5)
    order = np.lexsort((arr[:, 4], arr[:, 3], arr[:, 2], arr[:, 1], arr[:, 0]))
    return arr[order]


def main():
    ref.build()
    # 1. toy FASTA reads, both strands, l = 3 (11 rows)
    ents = both_strand_entries(TOY_READS)
    toy = {"reads": TOY_READS, "min_length": 3, "rows": ref.ref_overlaps(ents, 3)}
    assert len(toy["rows"]) == 11
    # 2. random tiny sets (small alphabets, injected substrings/duplicates, empty strings)
    tiny = tiny_cases(20261104, 300)
    with open(os.path.join(OUT, "tiny.json"), "w") as f:
        json.dump({"toy": toy, "tiny": tiny}, f, separators=(",", ":"))
    # 3. edge mix (duplicates, contained reads, short reads, periodic, palindromic)
    mids = {}
    for seed, l in ((3, 200), (4, 33), (5, 20)):
        reads = synth.edge_mix(seed, l)
        ents = both_strand_entries(reads)
        rows = ref.ref_overlaps(ents, l)
        mids["edge_s%d_l%d" % (seed, l)] = index_rows(rows, [i for i, _ in ents])
        print("edge", seed, l, len(ents), len(rows), "dups", len(rows) - len(set(rows)))
    # 4. the five BASELINE config shapes scaled down to what the reference does in seconds
    for cid, factor in ((1, 0.02), (2, 0.015), (3, 0.002), (4, 0.004), (5, 0.001)):
        sh = synth.CONFIGS[cid].scaled(factor)
        buf, off, names = synth.generate(sh)
        reads = [(n, buf[int(off[i]):int(off[i + 1])].tobytes().decode())
                 for i, n in enumerate(names)]
        ents = both_strand_entries(reads)
        rows = ref.ref_overlaps(ents, sh.min_length)
        mids["cfg%d_x%g" % (cid, factor)] = index_rows(rows, [i for i, _ in ents])
        print("cfg", cid, factor, sh.n_bases, len(ents), len(rows),
              "dups", len(rows) - len(set(rows)))
    np.savez_compressed(os.path.join(OUT, "mid.npz"), **mids)


if __name__ == "__main__":
    main()
