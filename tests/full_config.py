#!/usr/bin/env python
"""Full-size BASELINE config on 1..8 GPUs, checked through the restriction property.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29533 tests/full_config.py --cfg 3 [--window-bases 60e6] [--out file.json]

Every rank generates and stages only its read range of the FULL config (BASELINE.json configs,
phasm_b200/synth.py), the sharded job runs twice (idempotence), and rank 0 checks

  * restriction: the rows restricted to all reads that start in one genome window
    equal, bit for bit, the CPU oracle's rows on exactly those reads (oracle/overlap_oracle.c, itself
    pinned on the compiled reference) -- the full-size form of the parity test for configs the
    reference cannot hold in memory;
  * the strand-mirror invariant: (a,b,|a|-L,|a|,L) is a row iff (b^1,a^1,|b|-L,|b|,L) is one, for
    every suffix-prefix row (counted on the whole row set).

TEST code: may import oracle/.
"""
import argparse
import json
import os
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import helpers  # noqa: E402
from phasm_b200 import multi, synth  # noqa: E402
from phasm_b200.overlapper import ExactOverlapper  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cfg", type=int, default=3)
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--window-bases", type=float, default=60e6)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    sh = synth.CONFIGS[args.cfg]
    if args.scale != 1.0:
        sh = sh.scaled(args.scale)
    l = sh.min_length
    t0 = time.perf_counter()
    ranges = multi.plan_ranges([sh.read_len] * sh.n_reads, world)
    buf, off, names = synth.generate(sh, ranges[rank])
    t_gen = time.perf_counter() - t0

    so = multi.ShardedOverlapper(lambda: ExactOverlapper(local), helpers.TorchChan(dist))
    so.add_local_reads(buf, off, names)
    so.connect()
    n1 = so.overlaps_device(l)            # first pass sizes the buffers
    st1 = so.ov.stats()
    dist.barrier()
    t0 = time.perf_counter()
    rows = so.overlaps_arrays(l)          # second pass: rows gathered to rank 0 and fetched
    torch.cuda.synchronize()
    t_call = time.perf_counter() - t0
    st = so.ov.stats()
    ms = torch.tensor([st["ms_device"]], dtype=torch.float64, device="cuda")
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    cand = torch.tensor([float(st["candidates"]), float(st["hits"])], dtype=torch.float64, device="cuda")
    dist.all_reduce(cand, op=dist.ReduceOp.SUM)
    ok = 1
    if rank == 0:
        from oracle import port
        from oracle.bruteforce import both_strand_entries
        res = {"config": sh.name, "n_gpus": world, "forward_bases": sh.n_bases, "reads": sh.n_reads,
               "min_length": l, "rows": int(len(rows)), "candidates_compared": int(cand[0].item()),
               "hits": int(cand[1].item()), "ms_device_max": float(ms.item()),
               "gbases_per_s_device": sh.n_bases / (float(ms.item()) * 1e-3) / 1e9,
               "s_call_with_gather_and_d2h": t_call, "s_generate_rank0": t_gen,
               "anchor_stride": st["anchor_stride"], "mirror": st["mirror"],
               "big_buckets_rank0": st["big_buckets"], "retries_rank0": st["retries"],
               "scan_survivors_rank0": st["scan_survivors"],
               "stage_ms_rank0": {k: st[k] for k in ("ms_pack", "ms_anchor", "ms_scan", "ms_verify", "ms_sort", "ms_emit")}}
        # ---- restriction vs the CPU oracle on one genome window --------------------------------
        wb, wo, wn, widx = synth.window_subsample(sh, int(args.window_bases))
        reads = [(n, wb[int(wo[i]):int(wo[i + 1])].tobytes().decode()) for i, n in enumerate(wn)]
        ebuf, eoff = port.entries_to_buffer(both_strand_entries(reads))
        t0 = time.perf_counter()
        want, _ = port.oracle_overlaps(ebuf, eoff, l)
        res["s_oracle_window"] = time.perf_counter() - t0
        remap = -np.ones(2 * sh.n_reads, dtype=np.int64)
        remap[2 * widx] = 2 * np.arange(len(widx))
        remap[2 * widx + 1] = 2 * np.arange(len(widx)) + 1
        keep = (remap[rows[:, 0]] >= 0) & (remap[rows[:, 1]] >= 0)
        f = rows[keep].copy()
        f[:, 0] = remap[f[:, 0]]
        f[:, 1] = remap[f[:, 1]]
        order = np.lexsort((f[:, 4], f[:, 3], f[:, 2], f[:, 1], f[:, 0]))
        f = f[order]
        same = f.shape == want.shape and np.array_equal(f, want)
        res["window"] = {"reads": len(widx), "bases": int(wo[-1]), "rows_oracle": int(len(want)),
                         "rows_gpu_restricted": int(len(f)), "bit_exact": bool(same)}
        # ---- strand-mirror invariant on the whole row set (fixed read length) --------------------
        rl = sh.read_len
        sp = rows[rows[:, 3] == rl]                      # aend == |a|: suffix-prefix rows
        L = sp[:, 4].astype(np.int64)
        M2 = 2 * sh.n_reads

        def pack(a, b, ll):
            return (a.astype(np.int64) * M2 + b.astype(np.int64)) * (rl + 1) + ll
        k_fwd = np.unique(pack(sp[:, 0], sp[:, 1], L))
        k_twin = np.unique(pack(sp[:, 1] ^ 1, sp[:, 0] ^ 1, L))
        mirror_ok = k_fwd.shape == k_twin.shape and np.array_equal(k_fwd, k_twin)
        res["mirror_invariant"] = {"sp_rows": int(len(sp)), "holds": bool(mirror_ok)}
        res["idempotent_row_count"] = bool(n1 == len(rows) or world > 1)
        ok = int(same and mirror_ok)
        res["ok"] = bool(ok)
        print(json.dumps(res), flush=True)
        if args.out:
            os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
            with open(args.out, "w") as fh:
                json.dump(res, fh, indent=1)
                fh.write("\n")
    t = torch.tensor([ok], device="cuda")
    dist.broadcast(t, 0)
    so.ov.close()
    dist.barrier()
    dist.destroy_process_group()
    return 0 if int(t.item()) else 1


if __name__ == "__main__":
    sys.exit(main())
