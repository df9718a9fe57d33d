"""Run the UNMODIFIED reference (oracle/_ref, T0) on BASELINE config 1 IN FULL (200 Mbases, 40 000
entries, -l 1000) and record wall time, peak RSS and a digest of its rows.

BASELINE.md section 3.4 / SURVEY.md 8(d) promise this run; the digest lets the GPU tests compare the
CUDA path with the real reference at a full BASELINE size (``tests/golden/config1_full.json``,
checked by ``tests/test_gpu_parity.py::test_config1_full_equals_the_reference_digest``), and pins
the C restatement (T2) to T0 at that size as well.

Usage (build container, where /root/reference is mounted):
    TMPDIR=/tmp python tests/golden/make_config1_full.py [--config 1] [--out tests/golden/config1_full.json]

Follows the reference's caller ``phasm/cli/assembler.py:31-42``: ``add_sequence(name+'+', seq)``,
``add_sequence(name+'-', revcomp(seq))``, ``overlaps(l)``.
"""
from __future__ import annotations

import argparse
import json
import os
import resource
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import port, ref  # noqa: E402
from oracle.bruteforce import both_strand_entries  # noqa: E402
from phasm_b200 import synth  # noqa: E402
from helpers import rows_digest  # noqa: E402



def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, default=1)
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--out", default=os.path.join(ROOT, "tests", "golden", "config1_full.json"))
    ap.add_argument("--window-bases", type=float, default=0.0,
                    help="configs the reference cannot hold in memory: all reads of the FULL config that start in one "
                         "genome window of about this many forward bases (synth.window_subsample, the restriction "
                         "property of SURVEY.md 3.3) instead of the whole config")
    args = ap.parse_args()
    shape = synth.CONFIGS[args.config]
    if args.scale != 1.0:
        shape = shape.scaled(args.scale)
    if args.window_bases > 0:
        buf, off, names, _ = synth.window_subsample(shape, int(args.window_bases))
    else:
        buf, off, names = synth.generate(shape)
    reads = [(n, buf[int(off[i]):int(off[i + 1])].tobytes().decode()) for i, n in enumerate(names)]
    ents = both_strand_entries(reads)
    index = {eid: i for i, (eid, _) in enumerate(ents)}
    rec = {"shape": shape.name, "window_bases": int(args.window_bases), "seed": shape.seed, "reads": len(reads), "entries": len(ents),
           "forward_bases": int(off[-1]), "min_length": shape.min_length,
           "host": {"cores": os.cpu_count(), "threads_used": 1,
                    "tmpdir": os.environ.get("TMPDIR", "/tmp")},
           "reference": "oracle/_ref (unmodified /root/reference/src/{overlapper,phasm}.cpp, g++ -O3 -DNDEBUG)"}

    # T2 first (seconds), so that a killed T0 run still leaves something to compare
    ebuf, eoff = port.entries_to_buffer(ents)
    t0 = time.perf_counter()
    want, _ = port.oracle_overlaps(ebuf, eoff, shape.min_length)
    rec["t2_seconds"] = time.perf_counter() - t0
    rec["t2"] = rows_digest(want)
    print("T2:", rec["t2"], "%.1f s" % rec["t2_seconds"], flush=True)

    ov = ref.load().ExactOverlapper()
    t0 = time.perf_counter()
    for eid, seq in ents:
        ov.add_sequence(eid, seq)
    rec["t0_add_seconds"] = time.perf_counter() - t0
    t0 = time.perf_counter()
    res = ov.overlaps(int(shape.min_length))
    rec["t0_overlaps_seconds"] = time.perf_counter() - t0
    rec["t0_peak_rss_gb"] = resource.getrusage(resource.RUSAGE_SELF).ru_maxrss / 1e6
    got = np.empty((len(res), 5), dtype=np.uint32)
    for i, (a, b, s, e, bs, be) in enumerate(res):
        assert bs == 0
        got[i] = (index[a], index[b], s, e, be)
    rec["t0"] = rows_digest(got)
    rec["t0_gbases_per_s"] = rec["forward_bases"] / rec["t0_overlaps_seconds"] / 1e9
    rec["t0_overlaps_per_s"] = len(res) / rec["t0_overlaps_seconds"]
    rec["t2_equals_t0"] = rec["t0"] == rec["t2"]
    print(json.dumps(rec, indent=1))
    with open(args.out, "w") as f:
        json.dump(rec, f, indent=1)
        f.write("\n")


if __name__ == "__main__":
    main()
