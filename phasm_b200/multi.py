"""Read-range sharding of the overlapper over N GPUs of one box (one process per GPU).

SURVEY.md 8(e): the entry table is partitioned into contiguous A ranges balanced on bases; rank
r stages only its own reads, learns the lengths of all the others (so that every rank sees the
same entry table) and packs its range.  Everything that moves between the GPUs during a pass
(prefix table, filter slices, packed reads, hit keys, rows) moves inside the library over
peer-mapped HBM (``phasm_b200/csrc/k6_exchange.cuh``); this module is only the host-side
plumbing around it:

* ``plan_ranges``        contiguous, base-balanced read ranges
* ``TcpRendezvous``      a byte all-gather between the ranks over localhost TCP (standard library
                         only): the channel for the read lengths and the 128-byte NCCL id
* ``ShardedOverlapper``  builds the identical entry table on every rank from rank-local reads
* ``spawn_workers``      what ``phasm overlap --gpus N`` uses: N-1 child processes + this one

No torch here: any object with ``rank``, ``world`` and ``allgather_bytes(bytes) -> [bytes]`` works
as the channel (``bench.py`` and the tests pass an adapter over ``torch.distributed``, NCCL or
gloo, because torchrun has already set one up for them).

The reference has no counterpart (single process, single thread: SURVEY.md 2.1).
"""
from __future__ import annotations

import os
import socket
import struct
import subprocess
import sys
import time
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


# ---- byte channel between the ranks ---------------------------------------------------------------
def _send_blob(sock: socket.socket, blob: bytes) -> None:
    sock.sendall(struct.pack("<Q", len(blob)) + blob)


def _recv_exact(sock: socket.socket, n: int) -> bytes:
    parts = []
    while n:
        chunk = sock.recv(min(n, 1 << 20))
        if not chunk:
            raise ConnectionError("rendezvous peer closed the connection")
        parts.append(chunk)
        n -= len(chunk)
    return b"".join(parts)


def _recv_blob(sock: socket.socket) -> bytes:
    (n,) = struct.unpack("<Q", _recv_exact(sock, 8))
    return _recv_exact(sock, n)


class TcpRendezvous:
    """All-gather of byte strings between ``world`` processes of one host: rank 0 listens on
    ``addr:port``, the others connect and stay connected.  Only small control data travels here."""

    def __init__(self, rank: int, world: int, addr: str = "127.0.0.1", port: int = 29571,
                 timeout: float = 120.0):
        self.rank, self.world = rank, world
        self._peers: List[Optional[socket.socket]] = [None] * world
        self._sock: Optional[socket.socket] = None
        if world == 1:
            return
        if rank == 0:
            srv = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
            srv.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
            srv.bind((addr, port))
            srv.listen(world)
            srv.settimeout(timeout)
            try:
                for _ in range(world - 1):
                    conn, _ = srv.accept()
                    conn.settimeout(timeout)
                    (r,) = struct.unpack("<I", _recv_exact(conn, 4))
                    if not 0 < r < world or self._peers[r] is not None:
                        raise ConnectionError("rendezvous: unexpected rank %d" % r)
                    self._peers[r] = conn
            finally:
                srv.close()
        else:
            deadline = time.time() + timeout
            while True:
                s = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
                try:
                    s.connect((addr, port))
                    break
                except OSError:
                    s.close()
                    if time.time() > deadline:
                        raise
                    time.sleep(0.05)
            s.settimeout(timeout)
            s.sendall(struct.pack("<I", rank))
            self._sock = s

    def allgather_bytes(self, blob: bytes) -> List[bytes]:
        if self.world == 1:
            return [blob]
        if self.rank == 0:
            got = [blob] + [_recv_blob(self._peers[r]) for r in range(1, self.world)]
            packed = b"".join(struct.pack("<Q", len(b)) + b for b in got)
            for r in range(1, self.world):
                _send_blob(self._peers[r], packed)
            return got
        _send_blob(self._sock, blob)
        packed = _recv_blob(self._sock)
        out, pos = [], 0
        for _ in range(self.world):
            (n,) = struct.unpack_from("<Q", packed, pos)
            out.append(packed[pos + 8:pos + 8 + n])
            pos += 8 + n
        return out

    def barrier(self) -> None:
        self.allgather_bytes(b"")

    def close(self) -> None:
        for s in self._peers + [self._sock]:
            if s is not None:
                try:
                    s.close()
                except OSError:
                    pass
        self._peers = [None] * self.world
        self._sock = None


class ShardedOverlapper:
    """One rank's view of an N-GPU overlap job.

    ``make_overlapper`` builds the per-rank device object (``ExactOverlapper(local_rank)`` in
    production; a stand-in with the same methods in the CPU tests).  ``chan`` is the byte channel
    between the ranks (``TcpRendezvous`` or anything with the same three members).
    """

    def __init__(self, make_overlapper, chan):
        self.rank, self.world, self.chan = chan.rank, chan.world, chan
        self.ov = make_overlapper()
        self.ranges: Optional[List[Tuple[int, int]]] = None

    def add_local_reads(self, buf, offsets, names, both_strands: bool = True):
        """Every rank passes ITS reads (rank order = read order).  Builds the same entry table
        everywhere: remote reads are declared by length only."""
        offsets = np.asarray(offsets, dtype=np.uint64)
        local_len = np.diff(offsets).astype(np.uint32)
        all_len = [np.frombuffer(b, dtype=np.uint32) for b in self.chan.allgather_bytes(local_len.tobytes())]
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
        """NCCL bootstrap communicator of the library (id made on rank 0, shipped through the
        channel), upload, range check.  The peer-memory arenas are mapped by the first pass."""
        if self.world > 1:
            raw = type(self.ov).comm_unique_id() if self.rank == 0 else b""
            uid = self.chan.allgather_bytes(raw)[0]
            self.ov.comm_init(uid, self.rank, self.world)
        self.ov.upload()
        if self.world > 1:
            self.ov.comm_sync_ranges()

    def overlaps_device(self, min_length: int) -> int:
        """Device pass on every rank; every rank keeps the rows of its A range in its HBM."""
        return self.ov.overlaps_device(min_length)

    def overlaps_arrays(self, min_length: int) -> np.ndarray:
        """All rows on rank 0 (canonical order; pulled from the other ranks' HBM), an empty array
        elsewhere.  The ranks meet again before returning: nobody starts another pass or tears its
        arena down while rank 0 is still pulling."""
        self.ov.overlaps_device(min_length)
        rows = self.ov.fetch_rows()
        if self.world > 1:
            self.chan.allgather_bytes(b"")
        return rows


# ---- process fan-out for the CLI -----------------------------------------------------------------
def free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def spawn_workers(world: int, fasta: str, min_length: int, port: int) -> List[subprocess.Popen]:
    """Start ranks 1..world-1 as child processes (``python -m phasm_b200.worker``), the rank handed
    over in the environment.  The children never touch the output file."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    procs = []
    for r in range(1, world):
        env = dict(os.environ, PHASM_OVL_RANK=str(r), PHASM_OVL_WORLD=str(world), PHASM_OVL_PORT=str(port))
        env["PYTHONPATH"] = root + os.pathsep + env.get("PYTHONPATH", "")
        procs.append(subprocess.Popen([sys.executable, "-m", "phasm_b200.worker", fasta, str(int(min_length))],
                                      env=env, stdout=subprocess.DEVNULL))
    return procs
