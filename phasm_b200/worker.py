"""One rank of ``phasm overlap --gpus N`` (one process per GPU).

Rank 0 is the ``phasm`` process itself (it writes the GFA2 file); ranks 1..N-1 are children
started by ``phasm_b200.multi.spawn_workers`` as ``python -m phasm_b200.worker FASTA MIN_LENGTH``
with ``PHASM_OVL_RANK / PHASM_OVL_WORLD / PHASM_OVL_PORT`` in the environment.  Every rank reads
the same FASTA file and stages only its record range (``ovl_add_fasta_shard``); the 128-byte NCCL
id travels over a localhost TCP rendezvous; everything else moves between the GPUs inside the
library.  Replaces, for N GPUs, the body of ``overlap()`` in
/root/reference/phasm/cli/assembler.py:29-50 (the reference is single-process).
"""
from __future__ import annotations

import os
import sys

os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")   # 9 busy streams per pass (ovl_create)

from .multi import TcpRendezvous
from .overlapper import ExactOverlapper


def run_rank(rank: int, world: int, port: int, fasta: str, min_length: int, out=None,
             device: int | None = None) -> dict:
    """Run this rank's share; rank 0 writes header, S lines and E lines to ``out`` (a file object
    with ``fileno()``).  Returns the rank's stats dict."""
    chan = TcpRendezvous(rank, world, port=port)
    ov = ExactOverlapper(rank if device is None else device)
    try:
        ov.add_fasta_shard(fasta, rank, world, True)
        raw = ExactOverlapper.comm_unique_id() if rank == 0 else b""
        uid = chan.allgather_bytes(raw)[0]
        ov.comm_init(uid, rank, world)
        ov.upload()
        ov.comm_sync_ranges()
        ov.overlaps_device(min_length)
        st = ov.stats()
        if rank == 0 and world > 1:
            ov.comm_gather_rows_device()    # the peers' rows into rank 0's HBM: the last time a peer is touched
        # nobody tears its arena down while a peer may still touch it.  The barrier sits BEFORE the
        # file is written: formatting a large GFA2 file can outlast the rendezvous' socket timeout,
        # and the workers have nothing left to contribute once their rows are on rank 0.
        chan.barrier()
        if rank == 0 and out is not None:
            out.flush()
            ov.write_gfa(out.fileno(), True, True, True)
        return st
    finally:
        ov.close()
        chan.close()


def main(argv=None) -> int:
    argv = sys.argv[1:] if argv is None else argv
    fasta, min_length = argv[0], int(argv[1])
    rank = int(os.environ["PHASM_OVL_RANK"])
    world = int(os.environ["PHASM_OVL_WORLD"])
    port = int(os.environ["PHASM_OVL_PORT"])
    run_rank(rank, world, port, fasta, min_length)
    return 0


if __name__ == "__main__":
    sys.exit(main())
