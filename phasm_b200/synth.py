"""Synthetic error-free read sets of the BASELINE.json config shapes.

numpy ``default_rng(seed)`` (PCG64), ``seed = 1000 + config_id``.  Base haplotype = iid uniform
ACGT of length G; every further haplotype = copy + substitutions at ``snp_rate`` of positions;
config 4 additionally copies ``n_dups`` segments of ``dup_len`` bases to random targets before
the haplotypes are derived.  Per haplotype ``G*cov/RL`` reads of exactly RL bases, start
~U[0, G-RL], strand ~Bernoulli(1/2).  Only the sequenced strand is stored; ``-`` entries are
derived by the overlapper (as phasm/cli/assembler.py:39-40 does).  Names are ``r%07d``.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Tuple

import numpy as np

_ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)
_COMP = np.zeros(256, dtype=np.uint8)
for _a, _b in zip(b"ACGT", b"TGCA"):
    _COMP[_a] = _b


@dataclass(frozen=True)
class Shape:
    name: str
    genome: int          # haplotype length G
    ploidy: int
    cov: int             # coverage per haplotype
    read_len: int
    min_length: int
    seed: int
    snp_rate: float = 1e-3
    n_dups: int = 0
    dup_len: int = 20000
    dup_div: float = 2e-3

    @property
    def n_reads(self) -> int:
        return self.ploidy * (self.genome * self.cov // self.read_len)

    @property
    def n_bases(self) -> int:
        return self.n_reads * self.read_len

    def scaled(self, factor: float, name: Optional[str] = None) -> "Shape":
        """Same shape on a genome ``factor`` times as long (reads/coverage shape unchanged)."""
        g = max(int(self.genome * factor), 2 * self.read_len)
        nd = int(round(self.n_dups * factor)) if self.n_dups else 0
        return Shape(name or f"{self.name}x{factor:g}", g, self.ploidy, self.cov, self.read_len,
                     self.min_length, self.seed, self.snp_rate, nd, self.dup_len, self.dup_div)


# BASELINE.json "configs", in order (config_id 1..5)
CONFIGS = {
    1: Shape("cfg1-diploid5Mb-20x10kb-l1000", 5_000_000, 2, 20, 10_000, 1000, 1001),
    2: Shape("cfg2-triploid4.6Mb-30x15kb-l1000", 4_600_000, 3, 30, 15_000, 1000, 1002),
    3: Shape("cfg3-tetraploid50Mb-40x20kb-l2000", 50_000_000, 4, 40, 20_000, 2000, 1003),
    4: Shape("cfg4-diploid20Mb-segdup-30x10kb-l500", 20_000_000, 2, 30, 10_000, 500, 1004,
             n_dups=200),
    5: Shape("cfg5-hexaploid100Mb-30x20kb-l1000", 100_000_000, 6, 30, 20_000, 1000, 1005),
}


def make_haplotypes(shape: Shape, rng: np.random.Generator) -> List[np.ndarray]:
    g = shape.genome
    base = _ACGT[rng.integers(0, 4, size=g, dtype=np.uint8)]
    for d in range(shape.n_dups):
        if g <= 2 * shape.dup_len:
            break
        src = int(rng.integers(0, g - shape.dup_len))
        dst = int(rng.integers(0, g - shape.dup_len))
        seg = base[src:src + shape.dup_len].copy()
        if d % 2 == 1:  # half of the copies diverged
            nm = int(shape.dup_len * shape.dup_div)
            pos = rng.integers(0, shape.dup_len, size=nm)
            seg[pos] = _ACGT[(np.searchsorted(_ACGT_SORTED, seg[pos]) + rng.integers(1, 4, size=nm)) % 4]
        base[dst:dst + shape.dup_len] = seg
    haps = [base]
    for _ in range(1, shape.ploidy):
        h = base.copy()
        nm = int(g * shape.snp_rate)
        pos = rng.integers(0, g, size=nm)
        h[pos] = _ACGT_BY_RANK[(np.searchsorted(_ACGT_SORTED, h[pos]) + rng.integers(1, 4, size=nm)) % 4]
        haps.append(h)
    return haps


_ACGT_SORTED = np.sort(_ACGT)          # A C G T
_ACGT_BY_RANK = _ACGT_SORTED


def generate(shape: Shape, read_range: Optional[Tuple[int, int]] = None):
    """Return ``(buf, offsets, names)``: uint8 ASCII of all (or ``read_range``) reads
    concatenated, uint64 offsets[R+1], and the read names.

    The full start/strand/haplotype draw is always made, so a ``read_range`` slice is the same
    reads the full call would produce (multi-GPU ranks generate only their shard).
    """
    rng = np.random.default_rng(shape.seed)
    haps = make_haplotypes(shape, rng)
    per_hap = shape.genome * shape.cov // shape.read_len
    rl = shape.read_len
    r_total = per_hap * shape.ploidy
    starts = rng.integers(0, shape.genome - rl + 1, size=r_total)
    strands = rng.integers(0, 2, size=r_total, dtype=np.uint8)
    order = rng.permutation(r_total)   # interleave haplotypes like a real run would
    lo, hi = (0, r_total) if read_range is None else read_range
    n = hi - lo
    buf = np.empty(n * rl, dtype=np.uint8)
    for j in range(n):
        i = int(order[lo + j])
        h = haps[i // per_hap]
        s = int(starts[i])
        seg = h[s:s + rl]
        if strands[i]:
            seg = _COMP[seg[::-1]]
        buf[j * rl:(j + 1) * rl] = seg
    offsets = (np.arange(n + 1, dtype=np.uint64) * np.uint64(rl))
    names = ["r%07d" % (lo + j) for j in range(n)]
    return buf, offsets, names


def window_subsample(shape: Shape, max_bases: int):
    """All reads (every haplotype) whose start lies in one genome window, sized to about
    ``max_bases`` forward bases: the bounded sample the CPU reference is timed on
    (BASELINE.md section 3.4).  Overlaps are decided per read pair, so the rows on this sample
    equal the full run's rows restricted to these reads (restriction property).
    Returns ``(buf, offsets, names, read_indices)``.
    """
    rng = np.random.default_rng(shape.seed)
    haps = make_haplotypes(shape, rng)
    per_hap = shape.genome * shape.cov // shape.read_len
    rl = shape.read_len
    r_total = per_hap * shape.ploidy
    starts = rng.integers(0, shape.genome - rl + 1, size=r_total)
    strands = rng.integers(0, 2, size=r_total, dtype=np.uint8)
    order = rng.permutation(r_total)
    frac = min(1.0, max_bases / float(r_total * rl))
    win = int(shape.genome * frac)
    w0 = (shape.genome - win) // 2
    pos_of = np.empty(r_total, dtype=np.int64)
    pos_of[order] = np.arange(r_total)
    sel = np.nonzero((starts >= w0) & (starts < w0 + win))[0]
    sel = sel[np.argsort(pos_of[sel])]         # keep FASTA order
    n = len(sel)
    buf = np.empty(n * rl, dtype=np.uint8)
    for j, i in enumerate(sel.tolist()):
        seg = haps[i // per_hap][int(starts[i]):int(starts[i]) + rl]
        if strands[i]:
            seg = _COMP[seg[::-1]]
        buf[j * rl:(j + 1) * rl] = seg
    offsets = (np.arange(n + 1, dtype=np.uint64) * np.uint64(rl))
    idx = pos_of[sel]
    names = ["r%07d" % int(x) for x in idx]
    return buf, offsets, names, idx


def edge_mix(seed: int, min_length: int, n_reads: int = 200, genome: int = 40000,
             max_len: int = 3000):
    """Edge-case parity set: variable lengths, exact duplicates, contained
    reads, reads shorter than l, periodic and reverse-palindromic reads.  Returns a list of
    ``(name, seq)`` forward reads.
    """
    rng = np.random.default_rng(seed)
    g = _ACGT[rng.integers(0, 4, size=genome, dtype=np.uint8)]
    g2 = g.copy()
    pos = rng.integers(0, genome, size=max(genome // 500, 1))
    g2[pos] = _ACGT[rng.integers(0, 4, size=len(pos))]
    reads = []

    def take(h, s, ln, strand):
        seg = h[s:s + ln]
        if strand:
            seg = _COMP[seg[::-1]]
        return seg.tobytes().decode("ascii")

    lo = max(min_length // 2, 1)
    for i in range(n_reads):
        ln = int(rng.integers(lo, max_len + 1))
        s = int(rng.integers(0, genome - ln + 1))
        reads.append(take(g if rng.random() < 0.5 else g2, s, ln, int(rng.integers(0, 2))))
    extra = []
    for i in range(max(n_reads // 20, 2)):
        r = reads[int(rng.integers(0, len(reads)))]
        extra.append(r)                                             # exact duplicate
        if len(r) > min_length + 2:
            a = int(rng.integers(0, len(r) - min_length))
            b = int(rng.integers(a + min_length, len(r) + 1))
            extra.append(r[a:b])                                    # contained read
            extra.append(r[len(r) - (b - a):])                      # contained suffix (dup row)
            extra.append(r[:b - a])                                 # contained prefix
        if min_length > 1:
            extra.append(r[:int(rng.integers(1, min_length))])      # shorter than l
    unit = "ACGTTGCA"[: int(rng.integers(2, 8))]
    extra.append((unit * (max_len // len(unit) + 1))[:max_len])     # periodic
    extra.append((unit * (max_len // len(unit) + 1))[:max_len // 2])
    extra.append(("A" * max_len))                                   # homopolymer
    extra.append(("A" * (max_len // 3)))
    half = take(g, 0, max(min_length, 64), 0)
    rc = half.encode().translate(bytes.maketrans(b"ACGT", b"TGCA"))[::-1].decode()
    extra.append(half + rc)                                         # reverse-palindrome (== its RC)
    reads.extend(extra)
    perm = rng.permutation(len(reads))
    return [("e%05d" % j, reads[int(i)]) for j, i in enumerate(perm)]
