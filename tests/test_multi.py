"""Multi-GPU path: host plumbing on CPU (gloo, world_size 2) and, on a box with >= 2 GPUs, the
whole sharded job against the oracle (tests/mg_check.py under torchrun)."""
import os
import subprocess
import sys

import numpy as np
import pytest

import helpers

ROOT = helpers.ROOT


# ---- CPU: range planning ----------------------------------------------------------------------
def test_plan_ranges_tiles_and_balances():
    from phasm_b200.multi import plan_ranges
    rng = np.random.default_rng(3)
    for world in (1, 2, 3, 4, 8):
        for lens in ([15000] * 1000, rng.integers(1, 30000, size=777).tolist(), [5], [], [7] * 3):
            rs = plan_ranges(lens, world)
            assert len(rs) == world and rs[0][0] == 0 and rs[-1][1] == len(lens)
            assert all(rs[i][1] == rs[i + 1][0] for i in range(world - 1))
            assert all(lo <= hi for lo, hi in rs)
            if len(lens) >= 100 * world:
                tot = sum(lens)
                for lo, hi in rs:
                    assert abs(sum(lens[lo:hi]) - tot / world) <= max(lens) + 1
    with pytest.raises(ValueError):
        plan_ranges([1, 2], 0)


# ---- CPU: gloo world_size 2, stand-in device object ---------------------------------------------
_WORKER = r'''
import os, sys
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, os.path.join(sys.argv[1], "tests"))
import numpy as np, torch, torch.distributed as dist
import helpers
from phasm_b200 import multi

class FakeOverlapper:
    """Records what the sharding layer asks of the device object."""
    uid = bytes(range(128))
    def __init__(self): self.calls = []; self.lengths = []; self.local = []
    def add_reads_bulk(self, buf, off, names, both=True):
        ln = np.diff(np.asarray(off, dtype=np.uint64)).astype(np.uint32)
        self.local = [len(self.lengths), len(self.lengths) + len(ln)]
        self.lengths += ln.tolist(); self.calls.append(("bulk", len(ln), bool(both)))
    def add_remote(self, ln, both=True):
        self.lengths += list(map(int, ln)); self.calls.append(("remote", len(ln), bool(both)))
    @staticmethod
    def comm_unique_id(): return FakeOverlapper.uid
    def comm_init(self, uid, rank, world): self.calls.append(("init", bytes(uid), rank, world))
    def upload(self): self.calls.append(("upload",))
    def comm_sync_ranges(self): self.calls.append(("sync",))

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")
all_lens = [100, 2500, 30, 777, 1200, 64, 9000]
ranges = multi.plan_ranges(all_lens, world)
lo, hi = ranges[rank]
mine = all_lens[lo:hi]
off = np.concatenate(([0], np.cumsum(mine))).astype(np.uint64)
buf = np.full(int(off[-1]), 65, dtype=np.uint8)
so = multi.ShardedOverlapper(FakeOverlapper, helpers.TorchChan(dist))
got = so.add_local_reads(buf, off, ["r%d" % i for i in range(lo, hi)])
assert got == ranges, (got, ranges)
assert so.ov.lengths == all_lens, so.ov.lengths          # same entry table on every rank
assert so.ov.local == [lo, hi]
so.connect()
kinds = [c[0] for c in so.ov.calls]
assert kinds[-3:] == ["init", "upload", "sync"], kinds
init = so.ov.calls[-3]
assert init[1] == FakeOverlapper.uid and init[2] == rank and init[3] == world   # rank 0's id everywhere
dist.barrier(); dist.destroy_process_group()
print("worker %d ok" % rank)
'''


def test_sharded_plumbing_gloo_world2(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(_WORKER)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", "29517", str(script), ROOT]
    env = dict(os.environ, OMP_NUM_THREADS="1")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "worker 0 ok" in r.stdout and "worker 1 ok" in r.stdout


# ---- CPU: the torch-free byte channel the CLI uses ---------------------------------------------------
_RDV = r'''
import sys
sys.path.insert(0, sys.argv[1])
from phasm_b200.multi import TcpRendezvous
rank, world, port = int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
ch = TcpRendezvous(rank, world, port=port, timeout=60)
got = ch.allgather_bytes(bytes([rank]) * (rank * 1000 + 1))
assert [len(b) for b in got] == [r * 1000 + 1 for r in range(world)], [len(b) for b in got]
assert all(set(b) == {r} for r, b in enumerate(got))
uid = ch.allgather_bytes(b"U" * 128 if rank == 0 else b"")[0]      # how the NCCL id travels
assert uid == b"U" * 128
ch.barrier(); ch.close()
print("rdv %d ok" % rank)
'''


def test_placement_rule_keeps_exactly_one_of_every_twin_pair():
    """The multi-GPU placement rule (ovl_common.cuh: verifier_rank / mirror_items), evaluated on the
    host through the library: for every candidate item (suffix-prefix or containment) exactly one
    of {the item, its strand-mirror twin} is compared, both on the same rank, that rank holds one of
    the two reads, the read it does not hold enters as a forward strand, and rank r only ever needs
    the reads of ranks r+1 .. r+floor(N/2)."""
    import ctypes
    from phasm_b200 import _capi
    lib = _capi.load_library()
    rng = np.random.default_rng(7)
    for n_ranks in (1, 2, 3, 4, 5, 8):
        reads = 6 * n_ranks
        cuts = np.linspace(0, reads, n_ranks + 1).astype(np.uint32) * 2      # entry ranges, cut between reads
        rank_e = (ctypes.c_uint32 * (n_ranks + 1))(*[int(x) for x in cuts])
        lens = rng.integers(50, 200, size=reads)
        owner = lambda e: int(np.searchsorted(cuts, e, side="right") - 1)

        def rule(a, b, p):
            items, ver = ctypes.c_uint32(), ctypes.c_int()
            la, lb = int(lens[a >> 1]), int(lens[b >> 1])
            assert lib.ovl_debug_mirror_items(n_ranks, rank_e, a, la, b, lb, p, ctypes.byref(items), ctypes.byref(ver)) == 0
            return items.value, ver.value

        work = np.zeros(n_ranks)
        for _ in range(4000):
            a, b = (int(x) for x in rng.integers(0, 2 * reads, size=2))
            if a == b:
                continue
            la, lb = int(lens[a >> 1]), int(lens[b >> 1])
            p = int(rng.integers(0, la))
            rem = la - p
            mine, ver = rule(a, b, p)
            ra, rb = owner(a), owner(b)
            assert ver in (ra, rb)
            if ra != rb:
                d = (rb - ra) % n_ranks if ver == ra else (ra - rb) % n_ranks   # ver needs the rank d ahead of it
                assert 1 <= d <= n_ranks // 2
                if mine:
                    assert ((b if ver == ra else a) & 1) == 0      # the remote entry is a forward strand
            work[ver] += bin(mine).count("1")
            if rem <= lb:        # SP item: twin (b^1, lb-rem, a^1)
                t, tver = rule(b ^ 1, a ^ 1, lb - rem)
                assert tver == ver
                if (a, p, b) == (b ^ 1, lb - rem, a ^ 1):
                    assert mine & 2                                # its own twin: compares itself
                else:
                    assert bool(mine & 2) != bool(t & 2)
            if rem >= lb:        # CONT item: twin (a^1, la-p-lb, b^1)
                t, tver = rule(a ^ 1, b ^ 1, la - p - lb)
                assert tver == ver
                assert bool(mine & 1) != bool(t & 1)
        if n_ranks > 1:
            assert work.min() > 0.6 * work.mean(), work           # the compares spread over the ranks


def test_tcp_rendezvous_allgather_world3(tmp_path):
    from phasm_b200.multi import free_port
    script = tmp_path / "rdv.py"
    script.write_text(_RDV)
    port = free_port()
    procs = [subprocess.Popen([sys.executable, str(script), ROOT, str(r), "3", str(port)],
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for r in range(3)]
    outs = [p.communicate(timeout=120)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert all("rdv %d ok" % r in outs[r] for r in range(3)), outs


def test_product_packages_do_not_import_torch():
    import re
    for pkg in ("phasm_b200", "phasm"):
        for dirpath, _, files in os.walk(os.path.join(ROOT, pkg)):
            for f in files:
                if f.endswith(".py"):
                    src = open(os.path.join(dirpath, f)).read()
                    assert not re.search(r"^\s*(import|from)\s+torch\b", src, re.M), os.path.join(dirpath, f)


# ---- GPU: N ranks vs the oracle -------------------------------------------------------------------
def _n_gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.gpu
@pytest.mark.parametrize("extra", [["--cfg", "2", "--scale", "0.03"], ["--cfg", "4", "--scale", "0.01"],
                                   ["--edge", "3"], ["--cfg", "1", "--scale", "0.02", "--mirror", "0"],
                                   ["--cfg", "2", "--scale", "0.02", "--small-caps", "1"],
                                   ["--cfg", "2", "--scale", "0.01", "--bad-base", "1"],
                                   ["--edge", "4", "--lmin", "33"], ["--edge", "5", "--lmin", "20"],
                                   ["--skew", "7", "--lmin", "200", "--small-caps", "1"]])
def test_sharded_job_matches_oracle(extra):
    n = _n_gpus()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    world = (8 if n >= 8 else 4 if n >= 4 else 2) if "--edge" not in extra else 2
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", "29519",
           os.path.join(ROOT, "tests", "mg_check.py")] + extra
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert " OK" in r.stdout


@pytest.mark.gpu
def test_cli_overlap_gpus2_writes_the_same_bytes(tmp_path):
    """`phasm overlap --gpus 2` (one process per GPU over the TCP rendezvous, no torch) against
    `--gpus 1`: header and S lines byte-identical in order, E lines identical as lines in order too
    (both routes emit canonical order)."""
    if _n_gpus() < 2:
        pytest.skip("needs >= 2 GPUs")
    from phasm_b200 import synth
    sh = synth.CONFIGS[2].scaled(0.01)
    buf, off, names = synth.generate(sh)
    fa = tmp_path / "reads.fasta"
    with open(fa, "w") as f:
        for i, n in enumerate(names):
            f.write(">%s\n%s\n" % (n, buf[int(off[i]):int(off[i + 1])].tobytes().decode()))
    outs = []
    for g in (1, 2):
        out = tmp_path / ("out%d.gfa" % g)
        r = subprocess.run([sys.executable, "-m", "phasm", "overlap", "-l", str(sh.min_length), "--gpus", str(g),
                            "--stats", "json", "-o", str(out), str(fa)], capture_output=True, text=True, timeout=600, cwd=ROOT)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        assert "phasm_overlap_stats" in r.stderr
        outs.append(open(out, "rb").read())
    assert outs[0] == outs[1] and outs[0].count(b"\nE\t") > 1000


@pytest.mark.gpu
def test_fasta_shard_ingest_builds_the_same_entry_table_on_every_rank(tmp_path):
    """ovl_add_fasta_shard: what every rank of `phasm overlap --gpus N` does with the same file.  No
    communicator is needed to check the ingest, so three 'ranks' are three objects on one GPU."""
    from phasm_b200.overlapper import ExactOverlapper
    rng = np.random.default_rng(5)
    lens = [int(x) for x in rng.integers(1, 4000, size=61)] + [0, 7]
    fa = tmp_path / "r.fasta"
    with open(fa, "w") as f:
        for i, n in enumerate(lens):
            seq = "".join("ACGT"[int(c)] for c in rng.integers(0, 4, size=n))
            f.write(">read%d some description\n" % i)
            for k in range(0, n, 70):
                f.write(seq[k:k + 70] + "\n")
    world = 3
    ref = ExactOverlapper()
    assert ref.add_fasta(str(fa)) == len(lens)
    ids, lengths = ref.ids(), ref.lengths().tolist()
    ref.close()
    covered = 0
    for r in range(world):
        ov = ExactOverlapper()
        n, lo, hi = ov.add_fasta_shard(str(fa), r, world)
        assert n == len(lens) and lo == covered and lo <= hi
        covered = hi
        assert ov.ids() == ids and ov.lengths().tolist() == lengths       # same table everywhere
        if hi > lo:   # the rank's share is about 1/world of the bases
            assert abs(sum(lens[lo:hi]) - sum(lens) / world) <= max(lens)
        ov.close()
    assert covered == len(lens)
