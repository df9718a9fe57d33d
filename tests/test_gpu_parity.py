"""GPU parity tests: the CUDA path, called through the C ABI (ctypes -> libphasm_ovl.so), against

* the golden vectors generated from the COMPILED REFERENCE (tests/golden, make_golden.py),
* the C restatement oracle (oracle/overlap_oracle.c) on fresh seeded inputs,
* the brute-force restatement (oracle/bruteforce.py) on tiny random sets,
* size-independent properties (mirror invariant, restriction property, idempotence).

Bit-exact: the row multiset (a, b, astart, aend, bstart=0, bend) must be identical after a
canonical sort, duplicates included.
"""
import random

import numpy as np
import pytest

import helpers
from helpers import revcomp, sort_rows

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ov():
    from phasm_b200.overlapper import ExactOverlapper
    o = ExactOverlapper()
    yield o
    o.close()


def run_entries(ov, entries, l):
    ov.clear()
    for eid, seq in entries:
        ov.add_sequence(eid, seq)
    return sorted(ov.overlaps(l))


def run_reads_idx(ov, reads, l):
    ov.clear()
    for name, seq in reads:
        ov.add_read(name, seq)
    return sort_rows(ov.overlaps_arrays(l))


# ---- K1: pack + reverse complement ---------------------------------------------------------
def unpack(packed: np.ndarray, n: int) -> str:
    codes = np.zeros(len(packed) * 4, dtype=np.uint8)
    for j in range(4):
        codes[j::4] = (packed >> (2 * j)) & 3
    return np.frombuffer(b"ACTG", dtype=np.uint8)[codes[:n]].tobytes().decode()


def test_pack_and_revcomp_roundtrip(ov):
    rnd = random.Random(7)
    reads = []
    for n in [0, 1, 2, 15, 16, 17, 31, 32, 33, 63, 64, 65, 127, 128, 129, 1000, 4097]:
        reads.append(("p%d" % n, "".join(rnd.choice("ACGT") for _ in range(n))))
    ov.clear()
    for name, seq in reads:
        ov.add_read(name, seq)
    ov.overlaps_arrays(1 << 30)  # nothing is long enough: pack only
    for i, (name, seq) in enumerate(reads):
        fw = ov.debug_packed(2 * i)
        rc = ov.debug_packed(2 * i + 1)
        assert len(fw) == (len(seq) + 63) // 64 * 16
        assert unpack(fw, len(seq)) == seq
        assert unpack(rc, len(seq)) == revcomp(seq)
        # zero padding up to the granule boundary (the scan relies on it)
        assert unpack(fw, len(fw) * 4)[len(seq):] == "A" * (len(fw) * 4 - len(seq))
        assert unpack(rc, len(rc) * 4)[len(seq):] == "A" * (len(rc) * 4 - len(seq))


def test_non_acgt_is_rejected(ov):
    for bad in ("ACGTNACGT", "acgtacgtac", "ACGT ACGT", "ACGU", "ACG\x00"):
        ov.clear()
        ov.add_sequence("x", "ACGTACGTACGT")
        ov.add_sequence("bad", bad)
        with pytest.raises(ValueError):
            ov.overlaps(3)
    ov.clear()
    ov.add_read("y", "ACGTTTGN")
    with pytest.raises(ValueError):
        ov.overlaps(3)


# ---- golden vectors from the compiled reference ---------------------------------------------
def test_golden_toy_fasta(ov):
    g = helpers.load_tiny()["toy"]
    ov.clear()
    for name, seq in g["reads"]:
        ov.add_read(name, seq)
    got = sorted(ov.overlaps(g["min_length"]))
    assert got == sorted(g["rows"])
    assert len(got) == 11
    assert ("read1+", "read2+", 3, 7, 0, 4) in got


def test_golden_tiny_cases(ov):
    cases = helpers.load_tiny()["tiny"]
    assert len(cases) >= 300
    for i, c in enumerate(cases):
        got = run_entries(ov, c["entries"], c["min_length"])
        assert got == sorted(c["rows"]), "tiny case %d (l=%d): %r" % (i, c["min_length"], c["entries"])


@pytest.mark.parametrize("key", sorted(helpers.MID_SPECS))
def test_golden_mid(ov, key):
    want = sort_rows(helpers.load_mid()[key])
    reads, l = helpers.mid_input(key)
    got = run_reads_idx(ov, reads, l)
    assert got.shape == want.shape
    assert np.array_equal(got, want)


@pytest.mark.parametrize("stride", [1, 2, 8, 32, 64])
def test_anchor_stride_does_not_change_rows(ov, stride):
    """Every anchor stride (dense canonical scan .. every 64th position) gives the same rows."""
    try:
        ov.set_option("anchor_stride", stride)
        for key in ("edge_s3_l200", "cfg4_x0.004", "cfg2_x0.015"):
            want = sort_rows(helpers.load_mid()[key])
            reads, l = helpers.mid_input(key)
            got = run_reads_idx(ov, reads, l)
            assert np.array_equal(got, want), (key, stride)
            assert ov.stats()["anchor_stride"] == min(stride, 64)
    finally:
        ov.set_option("anchor_stride", 0)


def test_serial_verifier_gives_the_same_rows(ov):
    try:
        ov.set_option("serial_verify", 1)
        for key in ("edge_s4_l33", "cfg3_x0.002", "cfg5_x0.001"):
            want = sort_rows(helpers.load_mid()[key])
            reads, l = helpers.mid_input(key)
            assert np.array_equal(run_reads_idx(ov, reads, l), want), key
    finally:
        ov.set_option("serial_verify", 0)


def test_strand_mirror_halving_gives_the_same_rows(ov):
    """Mirror mode (compare one candidate of each strand-mirror pair, emit both rows) vs comparing
    every candidate: identical rows; it is on only when every entry has its twin."""
    for key in sorted(helpers.MID_SPECS):
        want = sort_rows(helpers.load_mid()[key])
        reads, l = helpers.mid_input(key)
        got = run_reads_idx(ov, reads, l)
        st = ov.stats()
        assert st["mirror"] == 1 and np.array_equal(got, want), key
        assert st["verified"] <= st["candidates"]
        try:
            ov.set_option("mirror", 0)
            got0 = run_reads_idx(ov, reads, l)
            st0 = ov.stats()
        finally:
            ov.set_option("mirror", 1)
        assert st0["mirror"] == 0 and st0["verified"] == st0["candidates"]
        assert np.array_equal(got0, want), key
    # add_sequence entries carry no twin information: mirror mode must stay off
    toy = helpers.load_tiny()["toy"]
    from oracle.bruteforce import both_strand_entries
    assert run_entries(ov, both_strand_entries(toy["reads"]), toy["min_length"]) == sorted(toy["rows"])
    assert ov.stats()["mirror"] == 0
    # mixed: reads + a lone sequence
    ov.clear()
    for name, seq in toy["reads"]:
        ov.add_read(name, seq)
    ov.add_sequence("lone", "GGGGGGGACTG")
    ov.overlaps_arrays(3)
    assert ov.stats()["mirror"] == 0


def test_mirror_tiny_random_both_strands(ov):
    """Tiny random read sets on both strands (palindromes, duplicates, containments, periodic
    reads) through add_read -> mirror mode, against the brute-force restatement."""
    from oracle.bruteforce import both_strand_entries, overlaps_bruteforce
    rnd = random.Random(77)
    for case in range(150):
        alpha = rnd.choice(["A", "AT", "ACGT", "ACGT"])
        reads = []
        for i in range(rnd.randint(1, 6)):
            s = "".join(rnd.choice(alpha) for _ in range(rnd.randint(1, 16)))
            if reads and rnd.random() < 0.4:
                t = rnd.choice(reads)[1]
                i0 = rnd.randrange(len(t)); i1 = rnd.randint(i0, len(t))
                s = rnd.choice([t[i0:i1] + s, s + t[i0:i1], t, revcomp(t), s + revcomp(s)])
            reads.append(("r%d" % i, s))
        l = rnd.randint(0, 5)
        ents = both_strand_entries(reads)
        want = sorted(overlaps_bruteforce(ents, l))
        ov.clear()
        for name, seq in reads:
            ov.add_read(name, seq)
        got = sorted(ov.overlaps(l))
        assert ov.stats()["mirror"] == 1
        assert got == want, (case, reads, l)


def test_auto_stride_follows_min_length(ov):
    """Auto stride: a power of two <= min(64, l-k+1) (the bound that keeps the scan complete);
    below l = 33 only the dense scan is possible.  Inside the bound a cost model picks it."""
    reads = helpers.mid_input("edge_s5_l20")[0]
    for l, bound in ((1, 1), (20, 1), (32, 1), (33, 2), (35, 4), (47, 16), (63, 32), (94, 32),
                     (95, 64), (1000, 64)):
        run_reads_idx(ov, reads, l)
        s = ov.stats()["anchor_stride"]
        assert 1 <= s <= bound and s & (s - 1) == 0, (l, s)
        assert (s == 1) == (bound == 1), (l, s)
        try:
            ov.set_option("anchor_stride", bound)   # an explicit value is an upper bound that is taken
            run_reads_idx(ov, reads, l)
            assert ov.stats()["anchor_stride"] == bound, l
        finally:
            ov.set_option("anchor_stride", 0)


# ---- oracle on fresh inputs -----------------------------------------------------------------
def test_bruteforce_random_tiny(ov):
    from oracle.bruteforce import both_strand_entries, overlaps_bruteforce
    rnd = random.Random(99)
    for it in range(150):
        alpha = rnd.choice(["A", "AC", "ACG", "ACGT"])
        ents = []
        for i in range(rnd.randint(1, 8)):
            s = "".join(rnd.choice(alpha) for _ in range(rnd.randint(0, 40)))
            if ents and rnd.random() < 0.4:
                t = rnd.choice(ents)[1]
                if t:
                    i0 = rnd.randrange(len(t))
                    i1 = rnd.randint(i0, len(t))
                    s = rnd.choice([t[i0:i1] + s, s + t[i0:i1], t[i0:i1], t])
            ents.append(("e%d" % i, s))
        if rnd.random() < 0.5:
            ents = both_strand_entries(ents)
        l = rnd.choice([0, 1, 2, 3, 5, 8, 13, 31, 32, 33, 34])
        assert run_entries(ov, ents, l) == overlaps_bruteforce(ents, l), (it, l, ents)


@pytest.mark.parametrize("cid,factor", [(1, 0.05), (2, 0.04), (4, 0.02)])
def test_oracle_port_config_shapes(ov, cid, factor):
    from oracle import port
    from phasm_b200 import synth
    sh = synth.CONFIGS[cid].scaled(factor)
    sh = synth.Shape(sh.name, sh.genome, sh.ploidy, sh.cov, sh.read_len, sh.min_length,
                     sh.seed + 77, sh.snp_rate, sh.n_dups, sh.dup_len, sh.dup_div)
    buf, off, names = synth.generate(sh)
    reads = helpers.reads_of(buf, off, names)
    from oracle.bruteforce import both_strand_entries
    ebuf, eoff = port.entries_to_buffer(both_strand_entries(reads))
    want, ost = port.oracle_overlaps(ebuf, eoff, sh.min_length)
    ov.clear()
    ov.set_collect_stats(True)
    ov.set_option("mirror", 0)  # the SURVEY 8(d) accounting is defined on ALL canonical candidates
    ov.add_reads_bulk(buf, off, names)
    try:
        got = sort_rows(ov.overlaps_arrays(sh.min_length))
        st = ov.stats()
        ov.set_option("mirror", 1)
        got_m = sort_rows(ov.overlaps_arrays(sh.min_length))
        st_m = ov.stats()
    finally:
        ov.set_option("mirror", 1)
        ov.set_collect_stats(False)
    assert np.array_equal(got, sort_rows(want))
    assert np.array_equal(got_m, sort_rows(want))
    # mirror mode compares about half of the candidates / operand bytes
    assert st_m["mirror"] == 1 and st_m["candidates"] < 0.6 * st["candidates"]
    assert st_m["operand_bytes"] < 0.6 * st["operand_bytes"]
    assert st["rows"] == ost["rows"] and st["hits"] == ost["hits"]
    assert st["anchors"] == ost["anchors"]
    assert st["scan_positions"] == ost["positions"]
    if sh.min_length >= 32:
        # same canonical candidate definition; self candidates are dropped
        # before the comparison on the device and never counted by the oracle
        assert st["candidates"] >= ost["candidates"]
        assert st["compared_bases"] == ost["compared_bases"]
        assert st["operand_bytes"] == ost["operand_bytes"]


def test_variable_lengths_and_short_l(ov):
    from oracle import port
    from oracle.bruteforce import both_strand_entries
    from phasm_b200 import synth
    for seed, l in ((11, 64), (12, 17), (13, 5)):
        reads = synth.edge_mix(seed, l, n_reads=120, genome=20000, max_len=1500)
        ebuf, eoff = port.entries_to_buffer(both_strand_entries(reads))
        want, _ = port.oracle_overlaps(ebuf, eoff, l)
        got = run_reads_idx(ov, reads, l)
        assert np.array_equal(got, sort_rows(want)), (seed, l)


# ---- properties -----------------------------------------------------------------------------
def test_mirror_invariant_and_idempotence(ov):
    from phasm_b200 import synth
    sh = synth.CONFIGS[1].scaled(0.03)
    buf, off, names = synth.generate(sh)
    ov.clear()
    ov.add_reads_bulk(buf, off, names)
    r1 = sort_rows(ov.overlaps_arrays(sh.min_length))
    r2 = sort_rows(ov.overlaps_arrays(sh.min_length))
    assert np.array_equal(r1, r2)                      # repeatable, index rebuilt per call
    lens = ov.lengths()
    rows = set(map(tuple, r1.tolist()))
    # (a,b,|a|-L,|a|,L) in R1  <=>  (b^,a^,|b|-L,|b|,L) in R1   
    for a, b, s, e, be in r1.tolist():
        if e == lens[a] and be == e - s:               # a suffix-prefix row
            L = be
            assert (b ^ 1, a ^ 1, int(lens[b]) - L, int(lens[b]), L) in rows
    assert len(r1) > 0


def test_restriction_property(ov):
    """overlaps(S') == overlaps(S) filtered to pairs inside S' (SURVEY.md 3.3)."""
    from phasm_b200 import synth
    sh = synth.CONFIGS[2].scaled(0.02)
    buf, off, names = synth.generate(sh)
    reads = helpers.reads_of(buf, off, names)
    full = run_reads_idx(ov, reads, sh.min_length)
    keep = np.arange(0, len(reads), 2)
    sub = run_reads_idx(ov, [reads[i] for i in keep], sh.min_length)
    remap = -np.ones(2 * len(reads), dtype=np.int64)
    for j, i in enumerate(keep):
        remap[2 * i] = 2 * j
        remap[2 * i + 1] = 2 * j + 1
    f = full[(remap[full[:, 0]] >= 0) & (remap[full[:, 1]] >= 0)].copy()
    f[:, 0] = remap[f[:, 0]]
    f[:, 1] = remap[f[:, 1]]
    assert np.array_equal(sort_rows(f), sub)


# ---- API behaviour of the reference binding ----------------------------------
def test_api_behaviour(ov):
    ov.clear()
    assert ov.overlaps(5) == []                         # the reference segfaults; we return []
    ov.add_sequence("only", "")
    assert ov.overlaps(0) == []
    ov.clear()
    ov.add_sequence("dup", "ACGTACGTAA")
    ov.add_sequence("dup", "ACGTAAGGGG")                 # ids are opaque: duplicates allowed
    ov.add_sequence(b"bytes-id", b"GGGGCC")              # bytes accepted
    rows = ov.overlaps(4)
    assert ("dup", "dup", 4, 10, 0, 6) in rows
    assert ("dup", "bytes-id", 6, 10, 0, 4) in rows
    with pytest.raises(TypeError):
        ov.overlaps(-1)
    with pytest.raises(TypeError):
        ov.overlaps(2.5)
    with pytest.raises(TypeError):
        ov.add_sequence("x", 5)
    # min_length 0 behaves as 1; can be called again after more add_sequence calls
    r0, r1 = ov.overlaps(0), ov.overlaps(1)
    assert r0 == r1
    ov.add_sequence("late", "CCAAAA")
    assert ("bytes-id", "late", 4, 6, 0, 2) in ov.overlaps(2)


def test_self_reverse_complement_pairs_are_reported(ov):
    # identity is by entry index: x+ vs x- ARE compared (overlapper.cpp:72-74)
    ov.clear()
    pal = "ACGTTGCATGCAACGT"                             # equals its own reverse complement
    assert revcomp(pal) == pal
    ov.add_read("pal", pal)
    rows = ov.overlaps(4)
    assert ("pal+", "pal-", 0, 16, 0, 16) in rows and ("pal-", "pal+", 0, 16, 0, 16) in rows
    assert rows.count(("pal+", "pal-", 0, 16, 0, 16)) == 2   # R1 + R2 duplicate row


# ---- hit ordering: warp buckets (default) and the block sort of big buckets -----------------------
def test_heavy_entries_are_sorted_by_the_block_kernel(ov):
    """An entry with more hits than one warp's bucket (k7_bucket.cuh BUCKET_MAX = 256) is ordered by
    k7_sort_big (one block per bucket, arbitrary length); rows must not change.  300 copies of one
    read give every entry ~600 hits, 700 copies ~1400 (more keys than the block has threads); a
    second, ordinary set afterwards must have no big bucket."""
    from oracle import port
    from oracle.bruteforce import both_strand_entries
    rng = random.Random(77)
    base = "".join(rng.choice("ACGT") for _ in range(300))
    other = "".join(rng.choice("ACGT") for _ in range(500))
    for copies in (300, 700):
        reads = [("d%03d" % i, base) for i in range(copies)] + [("x0", other), ("x1", other[100:] + base[:150])]
        ebuf, eoff = port.entries_to_buffer(both_strand_entries(reads))
        for l in (100, 300):
            want, _ = port.oracle_overlaps(ebuf, eoff, l)
            got = run_reads_idx(ov, reads, l)
            st = ov.stats()
            assert st["big_buckets"] > 0, st
            assert np.array_equal(got, want), (copies, l)
    key = sorted(helpers.MID_SPECS)[0]
    reads2, l2 = helpers.mid_input(key)
    assert np.array_equal(run_reads_idx(ov, reads2, l2), sort_rows(helpers.load_mid()[key]))
    assert ov.stats()["big_buckets"] == 0


def test_capacity_overflow_grows_and_repeats(ov):
    """A first candidate capacity far below what the input needs: the pass must notice the overflow
    on the device, grow, repeat and return the same rows (the retry route had no test)."""
    key = "cfg2_x0.015"
    reads, l = helpers.mid_input(key)
    want = sort_rows(helpers.load_mid()[key])
    ov.set_option("min_cand_cap", 64)
    ov.set_option("cand_per_entry", 1)
    try:
        got = run_reads_idx(ov, reads, l)
        st = ov.stats()
        assert st["retries"] > 0, st
        assert np.array_equal(got, want)
        got2 = run_reads_idx(ov, reads, l)          # capacities now fit: no retry, same rows
        assert np.array_equal(got2, want)
    finally:
        ov.set_option("min_cand_cap", 1 << 20)
        ov.set_option("cand_per_entry", 96)


def test_bucket_path_matches_serial_and_nonmirror_variants(ov):
    """The single-synchronisation pipeline (device-side counts) against the modes that still read
    counters on the host (serial verifier, accounting pass), mirror on and off."""
    from phasm_b200 import synth
    sh = synth.CONFIGS[4].scaled(0.01)
    buf, off, names = synth.generate(sh)
    reads = helpers.reads_of(buf, off, names)
    base = run_reads_idx(ov, reads, sh.min_length)
    assert len(base) > 0
    try:
        for opt, val in (("serial_verify", 1), ("mirror", 0)):
            ov.set_option(opt, val)
            assert np.array_equal(run_reads_idx(ov, reads, sh.min_length), base), opt
            ov.set_option(opt, 1 - val)
        ov.set_collect_stats(True)
        assert np.array_equal(run_reads_idx(ov, reads, sh.min_length), base)
    finally:
        ov.set_collect_stats(False)
        ov.set_option("serial_verify", 0)
        ov.set_option("mirror", 1)


def test_tuning_options_do_not_change_rows(ov):
    """Kernel variants behind ovl_set_option: bulk-async filter ring, candidate grouping before the
    verifier (the multi-GPU default), K2 / K1 grid sizes."""
    from phasm_b200 import synth
    sh = synth.CONFIGS[2].scaled(0.02)
    buf, off, names = synth.generate(sh)
    reads = helpers.reads_of(buf, off, names)
    base = run_reads_idx(ov, reads, sh.min_length)
    assert len(base) > 1000
    for opt, val, back in (("scan_bulk", 1, 0), ("group_cands", 1, -1), ("k2_blocks_per_sm", 2, 0),
                           ("pack_blocks_per_sm", 5, 16)):
        ov.set_option(opt, val)
        try:
            assert np.array_equal(run_reads_idx(ov, reads, sh.min_length), base), opt
        finally:
            ov.set_option(opt, back)
    ov.set_option("anchor_stride", 64)
    ov.set_option("scan_bulk", 1)
    try:
        assert np.array_equal(run_reads_idx(ov, reads, sh.min_length), base), "bulk stride 64"
    finally:
        ov.set_option("anchor_stride", 0)
        ov.set_option("scan_bulk", 0)


def test_config1_full_equals_the_reference_digest(ov):
    """BASELINE config 1 IN FULL (200 Mbases, 40 000 entries, -l 1000) against the UNMODIFIED reference:
    tests/golden/config1_full.json holds the row count and the sha256 of the sorted rows of the compiled
    reference's run on exactly this input (18 min, 14.5 GB on one host core;
    tests/golden/make_config1_full.py)."""
    import json
    import os
    from phasm_b200 import synth
    with open(os.path.join(helpers.GOLDEN, "config1_full.json")) as f:
        gold = json.load(f)
    sh = synth.CONFIGS[1]
    assert sh.name == gold["shape"] and sh.seed == gold["seed"]
    buf, off, names = synth.generate(sh)
    assert int(off[-1]) == gold["forward_bases"] and len(names) == gold["reads"]
    ov.clear()
    ov.add_reads_bulk(buf, off, names)
    got = ov.overlaps_arrays(sh.min_length, copy=False)
    assert helpers.rows_digest(got) == gold["t0"]
    ov.clear()


@pytest.mark.parametrize("cfg", [3, 4, 5])
def test_config_window_equals_the_reference_digest(ov, cfg):
    """A ~100 Mbase genome window of the FULL configs 3, 4 and 5 against the UNMODIFIED reference:
    tests/golden/config<k>_window.json holds the row count and the sha256 of the sorted rows of the
    compiled reference's run on exactly this input (about 10 min and 7-8 GB of one host core each;
    tests/golden/make_config1_full.py --window-bases 100e6).  Config 4 is the repeat-rich one
    (segmental duplications, -l 500: the candidate explosion case)."""
    import json
    import os
    from phasm_b200 import synth
    with open(os.path.join(helpers.GOLDEN, "config%d_window.json" % cfg)) as f:
        gold = json.load(f)
    sh = synth.CONFIGS[cfg]
    assert sh.name == gold["shape"] and sh.seed == gold["seed"]
    buf, off, names, _ = synth.window_subsample(sh, int(gold["window_bases"]))
    assert int(off[-1]) == gold["forward_bases"] and len(names) == gold["reads"]
    ov.clear()
    ov.add_reads_bulk(buf, off, names)
    got = ov.overlaps_arrays(sh.min_length, copy=False)
    assert len(got) == gold["t0"]["rows"]
    assert helpers.rows_digest(got) == gold["t0"]
    ov.clear()
