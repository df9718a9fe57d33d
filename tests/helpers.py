"""Shared test helpers (TEST code: may import oracle/)."""
import json
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")

_RC = bytes.maketrans(b"ACGT", b"TGCA")


def revcomp(s: str) -> str:
    return s.encode("ascii").translate(_RC)[::-1].decode("ascii")


def load_tiny():
    with open(os.path.join(GOLDEN, "tiny.json")) as f:
        d = json.load(f)
    for c in d["tiny"]:
        c["entries"] = [tuple(e) for e in c["entries"]]
        c["rows"] = [tuple(r) for r in c["rows"]]
    d["toy"]["reads"] = [tuple(r) for r in d["toy"]["reads"]]
    d["toy"]["rows"] = [tuple(r) for r in d["toy"]["rows"]]
    return d


def load_mid():
    return np.load(os.path.join(GOLDEN, "mid.npz"))


def sort_rows(arr: np.ndarray) -> np.ndarray:
    arr = np.asarray(arr, dtype=np.uint32).reshape(-1, 5)
    if len(arr) == 0:
        return arr
    order = np.lexsort((arr[:, 4], arr[:, 3], arr[:, 2], arr[:, 1], arr[:, 0]))
    return arr[order]


def rows_digest(rows) -> dict:
    """Row count + sha256 of the canonically sorted (a, b, astart, aend, bend) uint32 rows."""
    import hashlib
    rows = sort_rows(rows)
    return {"rows": int(len(rows)), "sha256": hashlib.sha256(np.ascontiguousarray(rows).tobytes()).hexdigest()}


def edges_digest(edges) -> str:
    """sha256 of an edge table [(u, v, edge_len, overlap_len, rank)] sorted by (u, v), as int64."""
    import hashlib
    arr = np.asarray(sorted(tuple(int(x) for x in e) for e in edges), dtype=np.int64).reshape(-1, 5)
    return hashlib.sha256(np.ascontiguousarray(arr).tobytes()).hexdigest()


def first_seq_to_rank(edges):
    """Edges (u, v, w, ol, first_seq) sorted by (u, v) -> the same with first_seq replaced by the rank
    of the edge in networkx's g.edges() order: nodes in order of first appearance (as u or v of an
    add_edge call), each node's out-edges in order of first insertion."""
    by_seq = sorted(edges, key=lambda e: e[4])
    node_order = {}
    for u, v, _, _, _ in by_seq:
        node_order.setdefault(u, len(node_order))
        node_order.setdefault(v, len(node_order))
    ranked = sorted(edges, key=lambda e: (node_order[e[0]], e[4]))
    rank = {(e[0], e[1]): r for r, e in enumerate(ranked)}
    return [(u, v, w, ol, rank[(u, v)]) for u, v, w, ol, _ in edges]


def reads_of(buf, off, names):
    return [(n, buf[int(off[i]):int(off[i + 1])].tobytes().decode()) for i, n in enumerate(names)]


def reads_to_buffer(reads):
    bs = [s.encode() for _, s in reads]
    off = np.zeros(len(bs) + 1, dtype=np.uint64)
    if bs:
        off[1:] = np.cumsum([len(b) for b in bs], dtype=np.uint64)
    buf = np.frombuffer(b"".join(bs), dtype=np.uint8) if bs else np.zeros(0, np.uint8)
    return buf, off, [n for n, _ in reads]


MID_SPECS = {
    # key in mid.npz -> how to regenerate the input
    "edge_s3_l200": ("edge", 3, 200),
    "edge_s4_l33": ("edge", 4, 33),
    "edge_s5_l20": ("edge", 5, 20),
    "cfg1_x0.02": ("cfg", 1, 0.02),
    "cfg2_x0.015": ("cfg", 2, 0.015),
    "cfg3_x0.002": ("cfg", 3, 0.002),
    "cfg4_x0.004": ("cfg", 4, 0.004),
    "cfg5_x0.001": ("cfg", 5, 0.001),
}


def mid_input(key):
    """-> (reads [(name, seq)], min_length) of a mid.npz golden vector."""
    from phasm_b200 import synth
    kind, x, y = MID_SPECS[key]
    if kind == "edge":
        return synth.edge_mix(x, y), y
    sh = synth.CONFIGS[x].scaled(y)
    buf, off, names = synth.generate(sh)
    return reads_of(buf, off, names), sh.min_length


def has_gpu() -> bool:
    import ctypes
    try:
        lib = ctypes.CDLL("libcudart.so.12")
    except OSError:
        return False
    n = ctypes.c_int(0)
    return lib.cudaGetDeviceCount(ctypes.byref(n)) == 0 and n.value > 0


class TorchChan:
    """Byte channel for phasm_b200.multi.ShardedOverlapper over an initialised torch.distributed
    process group (gloo on CPU, nccl under torchrun on a GPU box): the tests' stand-in for
    phasm_b200.multi.TcpRendezvous."""

    def __init__(self, dist):
        self.dist = dist
        self.rank, self.world = dist.get_rank(), dist.get_world_size()

    def allgather_bytes(self, blob: bytes):
        out = [None] * self.world
        self.dist.all_gather_object(out, bytes(blob))
        return out

    def barrier(self):
        self.dist.barrier()
