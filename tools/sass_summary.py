#!/usr/bin/env python
"""SASS evidence for the hot kernels: per kernel the register count, code size,
the memory / vote / shuffle / atomic opcode mix and whether any TMA / bulk-async / cp.async /
tensor-core opcode is present.  Usage: python tools/sass_summary.py [kernel regex ...] > profiles/<tag>_sass.txt"""
import collections
import re
import subprocess
import sys

LIB = "phasm_b200/_lib/libphasm_ovl.so"
pats = sys.argv[1:] or ["k1_pack", "k2_insert", "k3_filter_wide", "k3_resolve", "k4a_head", "k4b_tail", "k7_sort", "k7_emit"]
sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True).stdout
regs = {}
cur = None
for line in res.splitlines():
    m = re.search(r"Function (\S+):", line)
    if m:
        cur = m.group(1)
    m = re.search(r"REG:(\d+).*SHARED:(\d+)", line)
    if m and cur:
        regs[cur] = (int(m.group(1)), int(m.group(2)))
funcs = re.split(r"\n\s*Function : ", sass)[1:]
INTEREST = ("LDG", "STG", "LDS", "STS", "ATOM", "RED", "SHFL", "VOTE", "MATCH", "UTMA", "UBLKCP", "LDGSTS", "SYNCS",
            "UTC", "LDTM", "HMMA", "IMMA", "BAR", "SHF", "LOP3", "IMAD", "PRMT", "POPC", "BREV", "FLO")
print("# SASS summary of %s (sm_100a), cuobjdump -sass / -res-usage" % LIB)
for f in funcs:
    name = f.split("\n", 1)[0].strip()
    if not any(re.search(p, name) for p in pats):
        continue
    ops = collections.Counter()
    n = 0
    for line in f.splitlines():
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_.]+)", line)
        if m:
            n += 1
            ops[m.group(1)] += 1
    r = regs.get(name, ("?", "?"))
    print("\n== %s\n   registers %s, static shared %s B, %d SASS instructions" % (name, r[0], r[1], n))
    groups = collections.Counter()
    for op, c in ops.items():
        for k in INTEREST:
            if op.startswith(k):
                groups[op if k in ("LDG", "STG", "LDS", "STS", "ATOM", "RED") else k] += c
                break
    print("   " + ", ".join("%s x%d" % kv for kv in sorted(groups.items())))
    flags = [k for k in ("UTMALDG", "UTMASTG", "UBLKCP", "LDGSTS", "UTCMMA", "UTCHMMA", "LDTM", "HMMA") if any(o.startswith(k) for o in ops)]
    print("   TMA / bulk-async / cp.async / tensor-core opcodes: %s" % (", ".join(flags) if flags else "none"))
