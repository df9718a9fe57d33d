"""Read-range sharding of the overlapper over N GPUs of one box (one process per GPU).

SURVEY.md 8(e): the entry table is partitioned into contiguous A ranges balanced on bases; rank
r stages only its own reads, learns the lengths of all the others (so that every rank sees the
same entry table), packs its range, and the packed granules / hit keys are exchanged inside the
library over NCCL (``ovl_comm_*``).  This module is the host-side plumbing around that:

* ``plan_ranges``      contiguous, base-balanced read ranges
* ``ShardedOverlapper`` builds the identical entry table on every rank from rank-local reads,
  using a ``torch.distributed`` process group only for the small host-side exchanges (read
  lengths, the 128-byte NCCL id).  Works with the ``gloo`` backend too, which is what the CPU
  tests use (with a stand-in for the device object).

The reference has no counterpart (single process, single thread: SURVEY.md 2.1).
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import numpy as np


def plan_ranges(lengths: Sequence[int], world: int) -> List[Tuple[int, int]]:
    """Split reads ``0..R`` into ``world`` contiguous ranges with about equal base counts.

    Returns ``[(lo, hi)] * world``; ranges tile ``[0, R)`` in rank order; a range may be empty.
    """
    lengths = np.asarray(lengths, dtype=np.int64)
    r = len(lengths)
    if world < 1:
        raise ValueError("world must be >= 1")
    csum = np.concatenate(([0], np.cumsum(lengths)))
    total = int(csum[-1])
    cuts = [0]
    for k in range(1, world):
        target = total * k / world
        i = int(np.searchsorted(csum, target, side="left"))
        # pick the boundary whose prefix sum is closest to the target
        if i > 0 and abs(csum[i - 1] - target) <= abs(csum[min(i, r)] - target):
            i -= 1
        cuts.append(min(max(i, cuts[-1]), r))
    cuts.append(r)
    return [(cuts[k], cuts[k + 1]) for k in range(world)]


def _all_gather_lengths(local: np.ndarray, dist, device) -> List[np.ndarray]:
    """All-gather variable-length uint32 arrays through a process group."""
    import torch
    world = dist.get_world_size()
    n = torch.tensor([len(local)], dtype=torch.int64, device=device)
    ns = [torch.zeros(1, dtype=torch.int64, device=device) for _ in range(world)]
    dist.all_gather(ns, n)
    counts = [int(x.item()) for x in ns]
    cap = max(max(counts), 1)
    buf = torch.zeros(cap, dtype=torch.int64, device=device)
    if len(local):
        buf[:len(local)] = torch.from_numpy(local.astype(np.int64)).to(device)
    outs = [torch.zeros(cap, dtype=torch.int64, device=device) for _ in range(world)]
    dist.all_gather(outs, buf)
    return [o[:c].cpu().numpy().astype(np.uint32) for o, c in zip(outs, counts)]


class ShardedOverlapper:
    """One rank's view of an N-GPU overlap job.

    ``make_overlapper`` builds the per-rank device object (``ExactOverlapper(local_rank)`` in
    production; a stand-in with the same methods in the CPU tests).
    """

    def __init__(self, make_overlapper, rank: int, world: int, dist=None, device="cpu"):
        self.rank, self.world, self.dist, self.device = rank, world, dist, device
        self.ov = make_overlapper()
        self.ranges: Optional[List[Tuple[int, int]]] = None

    def add_local_reads(self, buf, offsets, names, both_strands: bool = True):
        """Every rank passes ITS reads (rank order = read order).  Builds the same entry table
        everywhere: remote reads are declared by length only."""
        offsets = np.asarray(offsets, dtype=np.uint64)
        local_len = np.diff(offsets).astype(np.uint32)
        if self.world > 1:
            all_len = _all_gather_lengths(local_len, self.dist, self.device)
        else:
            all_len = [local_len]
        lo = 0
        self.ranges = []
        for r, ln in enumerate(all_len):
            if r == self.rank:
                self.ov.add_reads_bulk(buf, offsets, names, both_strands)
            elif len(ln):
                self.ov.add_remote(ln, both_strands)
            self.ranges.append((lo, lo + len(ln)))
            lo += len(ln)
        return self.ranges

    def connect(self):
        """NCCL communicator of the library (id made on rank 0, broadcast through ``dist``),
        upload, range check."""
        if self.world > 1:
            import torch
            uid = torch.zeros(128, dtype=torch.uint8, device=self.device)
            if self.rank == 0:
                raw = type(self.ov).comm_unique_id()
                uid.copy_(torch.frombuffer(bytearray(raw), dtype=torch.uint8))
            self.dist.broadcast(uid, 0)
            self.ov.comm_init(bytes(uid.cpu().numpy().tobytes()), self.rank, self.world)
        self.ov.upload()
        if self.world > 1:
            self.ov.comm_sync_ranges()

    def overlaps_device(self, min_length: int) -> int:
        """Device pass on every rank + gather of the rows into rank 0's HBM."""
        n = self.ov.overlaps_device(min_length)
        return self.ov.comm_gather_rows_device() if self.world > 1 else n

    def overlaps_arrays(self, min_length: int) -> np.ndarray:
        """Rows on rank 0 (canonical order), an empty array elsewhere."""
        self.ov.overlaps_device(min_length)
        return self.ov.comm_gather_rows() if self.world > 1 else self.ov.fetch_rows()
