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
sys.path.insert(0, sys.argv[1])
import numpy as np, torch, torch.distributed as dist
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
so = multi.ShardedOverlapper(FakeOverlapper, rank, world, dist, "cpu")
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


# ---- GPU: N ranks vs the oracle -------------------------------------------------------------------
def _n_gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.gpu
@pytest.mark.parametrize("extra", [["--cfg", "2", "--scale", "0.03"], ["--cfg", "4", "--scale", "0.01"],
                                   ["--edge", "3"], ["--cfg", "1", "--scale", "0.02", "--mirror", "0"]])
def test_sharded_job_matches_oracle(extra):
    n = _n_gpus()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    world = 4 if n >= 4 and "--edge" not in extra else 2
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", "29519",
           os.path.join(ROOT, "tests", "mg_check.py")] + extra
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert " OK" in r.stdout
