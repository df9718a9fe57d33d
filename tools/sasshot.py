#!/usr/bin/env python
"""Hot SASS lines of one kernel of an ncu report: python tools/sasshot.py <rep> <kernel regex> [min share %]"""
import csv, subprocess, sys
rep, pat = sys.argv[1], sys.argv[2]
thr = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + pat],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = None
body = []
for r in rows:
    if "Instructions Executed" in r:
        if hdr is not None:
            break        # first kernel instance only
        hdr = r
        continue
    if hdr and len(r) == len(hdr):
        body.append(r)
ia, isrc, ist = hdr.index("Instructions Executed"), hdr.index("Source"), hdr.index("Warp Stall Sampling (All Samples)")
tot = sum(int(r[ia]) for r in body)
tots = sum(int(r[ist]) for r in body)
print("total warp-inst %d, stall samples %d, sass lines %d" % (tot, tots, len(body)))
for i, r in enumerate(body):
    n, s = int(r[ia]), int(r[ist])
    if 100.0 * n / max(tot, 1) >= thr or 100.0 * s / max(tots, 1) >= thr:
        print("%5d %10d %5.1f%% %7d %5.1f%%  %s" % (i, n, 100.0 * n / tot, s, 100.0 * s / max(tots, 1), r[isrc].strip()[:100]))
