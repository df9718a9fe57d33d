import sys; sys.path.insert(0,'.')
from phasm_b200.overlapper import ExactOverlapper
from oracle.bruteforce import both_strand_entries, overlaps_bruteforce
ov=ExactOverlapper()
def run(reads,l,mirror):
    ov.clear(); ov.set_option("mirror",mirror)
    for n,s in reads: ov.add_read(n,s)
    return sorted(ov.overlaps(l)), ov.stats()
reads=[('r0', 'CCTTTCTACGGGTCGC'), ('r1', 'GCCGCGTGG'), ('r2', 'GCGACCCGTAGAAAGG')]
for rs in (reads, [reads[0],reads[2]], [reads[0],reads[1]]):
  for l in (4,):
    want=sorted(overlaps_bruteforce(both_strand_entries(rs),l))
    g1,s1=run(rs,l,1); g0,s0=run(rs,l,0)
    print(len(rs), l, g1==want, g0==want, s1["candidates"], s1["hits"], s0["candidates"], s0["hits"])
    from collections import Counter
    print(" extra", Counter(g1)-Counter(want)); print(" missing", Counter(want)-Counter(g1))
