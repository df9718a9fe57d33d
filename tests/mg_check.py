#!/usr/bin/env python
"""Multi-GPU equivalence check (run under torchrun, one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29511 tests/mg_check.py [--cfg 2 --scale 0.05] [--mirror 1]

Every rank stages only its read range; rank 0 compares the gathered rows with the CPU oracle
(oracle/overlap_oracle.c) run on the whole read set.  TEST code: may import oracle/.
"""
import argparse
import os
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import helpers  # noqa: E402
from phasm_b200 import multi, synth  # noqa: E402
from phasm_b200.overlapper import ExactOverlapper  # noqa: E402


def skewed_reads(seed: int):
    """A read set whose candidates pile up on the FIRST read range: its reads are made of copies of
    one 300-base repeat R, and every other read starts with R, so every occurrence of R in the first
    range is a candidate against every other read while the other ranges attract next to nothing
    (the grow-and-repeat route when only one rank's inbound lists overflow).  A tiling of
    a random genome at the end provides real overlaps."""
    rng = np.random.default_rng(seed)
    acgt = np.frombuffer(b"ACGT", dtype=np.uint8)

    def rnd(n):
        return acgt[rng.integers(0, 4, n)].tobytes().decode()
    rep = rnd(300)
    reads = [("rep%d" % i, "".join(rep + rnd(75) for _ in range(4))) for i in range(30)]
    reads += [("pre%d" % i, rep + rnd(1200)) for i in range(60)]
    genome = rnd(24000)
    reads += [("til%d" % i, genome[600 * i:600 * i + 1500]) for i in range(38)]
    return reads


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cfg", type=int, default=2)
    ap.add_argument("--scale", type=float, default=0.05)
    ap.add_argument("--mirror", type=int, default=1)
    ap.add_argument("--edge", type=int, default=0, help="use synth.edge_mix(seed, l) instead of a config")
    ap.add_argument("--lmin", type=int, default=200, help="min_length of the --edge set")
    ap.add_argument("--skew", type=int, default=0, help="seed of a read set whose candidates concentrate on rank 0's reads")
    ap.add_argument("--small-caps", type=int, default=0, help="1: tiny first capacities -> the collective grow-and-repeat route")
    ap.add_argument("--bad-base", type=int, default=0, help="1: one non-ACGT byte on the LAST rank: every rank must raise")
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    if args.edge or args.skew:
        reads = skewed_reads(args.skew) if args.skew else synth.edge_mix(args.edge, args.lmin)
        l = args.lmin
        lens = [len(s) for _, s in reads]
        ranges = multi.plan_ranges(lens, world)
        lo, hi = ranges[rank]
        mine = reads[lo:hi]
        bs = [s.encode() for _, s in mine]
        off = np.zeros(len(bs) + 1, dtype=np.uint64)
        if bs:
            off[1:] = np.cumsum([len(b) for b in bs], dtype=np.uint64)
        buf = np.frombuffer(b"".join(bs), dtype=np.uint8) if bs else np.zeros(0, np.uint8)
        names = [n for n, _ in mine]
    else:
        sh = synth.CONFIGS[args.cfg].scaled(args.scale)
        l = sh.min_length
        ranges = multi.plan_ranges([sh.read_len] * sh.n_reads, world)
        buf, off, names = synth.generate(sh, ranges[rank])

    so = multi.ShardedOverlapper(lambda: ExactOverlapper(local), helpers.TorchChan(dist))
    so.ov.set_option("mirror", args.mirror)
    so.ov.set_option("peer_timeout_ms", 8000)
    if args.small_caps:
        so.ov.set_option("min_cand_cap", 64)
        so.ov.set_option("cand_per_entry", 1)
    if args.bad_base and rank == world - 1 and len(buf):
        buf = buf.copy()
        buf[len(buf) // 2] = ord("N")
    got_ranges = so.add_local_reads(buf, off, names)
    assert got_ranges == ranges, (got_ranges, ranges)
    so.connect()
    if args.bad_base:
        # a bad byte on one shard must raise on EVERY rank (status exchange), not hang the others
        try:
            so.overlaps_arrays(l)
            raised = False
        except ValueError as e:
            raised = "ACGT" in str(e)
        t = torch.tensor([int(raised)], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        ok = int(t.item())
        if rank == 0:
            print("mg_check world=%d bad-base: %s" % (world, "every rank raised OK" if ok else "MISMATCH"), flush=True)
        so.ov.close()
        dist.barrier()
        dist.destroy_process_group()
        return 0 if ok else 1
    rows = so.overlaps_arrays(l)
    rows2 = so.overlaps_arrays(l)          # repeatable
    st = so.ov.stats()
    ok = 1
    if rank == 0:
        from oracle import port
        from oracle.bruteforce import both_strand_entries
        if args.edge or args.skew:
            full = reads
        else:
            fb, fo, fn = synth.generate(sh)
            full = [(n, fb[int(fo[i]):int(fo[i + 1])].tobytes().decode()) for i, n in enumerate(fn)]
        ebuf, eoff = port.entries_to_buffer(both_strand_entries(full))
        want, _ = port.oracle_overlaps(ebuf, eoff, l)
        same = rows.shape == want.shape and np.array_equal(rows, want)   # already canonical order
        ok = int(same and np.array_equal(rows, rows2))
        if args.small_caps:
            ok = int(ok and st["retries"] == 0)    # second call: capacities fit
        print("mg_check world=%d mirror=%d rows=%d want=%d %s" % (world, st["mirror"], len(rows), len(want),
                                                                   "OK" if ok else "MISMATCH"), flush=True)
    print("rank %d: candidates %d hits %d sent %d received %d device %.3f ms" % (
        rank, st["candidates"], st["hits"], st["hits_sent"], st["hits_received"], st["ms_device"]), flush=True)
    t = torch.tensor([ok], device="cuda")
    dist.broadcast(t, 0)
    so.ov.close()
    dist.barrier()
    dist.destroy_process_group()
    return 0 if int(t.item()) else 1


if __name__ == "__main__":
    sys.exit(main())
