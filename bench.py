#!/usr/bin/env python
"""bench.py -- throughput of the B200 overlapper on BASELINE.json's metric.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

Metric  : Gbases scanned/s (forward input bases / time) and overlaps/s, all-vs-all exact overlap.
Workload: BASELINE.json configs[1] -- synthetic triploid 4.6 Mb genome, 30x per haplotype of
          15 kb error-free reads, -l 1000 (414 Mbases, 27 600 reads, 55 200 entries) on ONE
          B200.  At N GPUs the genome is N times as long (weak scaling: every rank owns the
          same number of reads, sharded by read range).
Step    : one pass of the hot path (pack -> anchors -> scan -> verify -> emit [-> gather to
          rank 0]) over the whole read set.
value   : inputs (ASCII reads + entry table) already resident in HBM; K steps bracketed by CUDA
          events on the stream the kernels are launched on, max over ranks.
e2e     : the same pass through the reference-facing call with HOST buffers: pinned host ASCII
          -> HBM, device path, rows -> host (ExactOverlapper.overlaps_arrays == ovl_overlaps),
          wall clock between device synchronisations, max over ranks.
roofline: the dominant kernel's algorithmic bytes / its CUDA-event
          duration inside the timed region, against MEASURED_PEAKS.json's HBM copy bandwidth.
cpu_baseline / --impl reference: the reference's own CPU overlapper (oracle/_ref, compiled
          from the reference's sources; else the C restatement) on a bounded genome-window
          sample of the same workload, on the host cores of this box.

torch is used for plumbing only (torch.distributed rendezvous / barrier / max-reduce and
torch.cuda.synchronize); the product path is libphasm_ovl.so through ctypes.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

from phasm_b200 import synth  # noqa: E402

METRIC = "Gbases scanned/s"
UNIT = "Gbases/s"
CONFIG_ID = 2  # BASELINE.json configs[1]


# ---------------------------------------------------------------------------------------------
def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic():
    """Per-kernel DRAM bytes per launch from the newest committed `ncu --set full` capture
    (profiles/<tag>_traffic.json, written by tools/profsum.py; same workload as this bench)."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "*_traffic.json")))
    if not files:
        return None, None
    try:
        with open(files[-1]) as f:
            k = json.load(f)["kernels"]
        return {n: v["dram_bytes_per_launch"] for n, v in k.items()}, os.path.basename(files[-1])
    except Exception:
        return None, None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu: int):
        self.gpu, self.proc = gpu, None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "20"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.05)
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
            out, _ = self.proc.communicate()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
                pw.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "power_w_max": max(pw),
                "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------
def cpu_reference_run(shape, sample_bases: int, repeats: int = 1):
    """Time the reference CPU overlapper on a genome-window sample of `shape`.
    Returns dict(value Gbases/s, seconds list, rows, bases, kind, cores, sample)."""
    from oracle import port, ref
    from oracle.bruteforce import both_strand_entries
    buf, off, names, _ = synth.window_subsample(shape, sample_bases)
    n_reads = len(names)
    bases = int(off[-1])
    reads = [(n, buf[int(off[i]):int(off[i + 1])].tobytes().decode()) for i, n in enumerate(names)]
    ents = both_strand_entries(reads)  # what phasm/cli/assembler.py:39-40 feeds the overlapper
    secs, rows = [], 0
    if ref.available():
        kind, cores = "reference", 1  # the reference is single-threaded
        mod = ref.load()
        for _ in range(repeats):
            ov = mod.ExactOverlapper()
            for eid, seq in ents:
                ov.add_sequence(eid, seq)
            t0 = time.perf_counter()
            res = ov.overlaps(int(shape.min_length))
            secs.append(time.perf_counter() - t0)
            rows = len(res)
            del res, ov
    else:
        kind, cores = "port", os.cpu_count() or 1
        ebuf, eoff = port.entries_to_buffer(ents)
        for _ in range(repeats):
            t0 = time.perf_counter()
            r, _ = port.oracle_overlaps(ebuf, eoff, shape.min_length, nthreads=cores)
            secs.append(time.perf_counter() - t0)
            rows = len(r)
    t = sum(secs) / len(secs)
    return {
        "value": bases / t / 1e9, "unit": UNIT, "cores": cores, "kind": kind,
        "host_cores": os.cpu_count(),
        "sample": "%d reads / %d forward bases (all reads starting in one genome window of %s), "
                  "overlaps(%d) wall time only, %.2f s per call"
                  % (n_reads, bases, shape.name, shape.min_length, t),
        "overlaps_per_s": rows / t, "rows": rows, "seconds": secs, "bases": bases,
    }


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    shape = workload_shape(args.gpus, args.scale)
    calls = max(args.steps + args.warmup, 1)
    per_step = min(max(150.0 / calls, 2.0), 30.0)            # seconds of CPU work per step
    sample = int(0.25e6 * per_step)                          # reference ~0.25 Mbases/s/core
    sample = max(sample, 4 * shape.read_len)
    for _ in range(args.warmup):
        cpu_reference_run(shape, sample, 1)
    r = cpu_reference_run(shape, sample, max(args.steps, 1))
    t = sum(r["seconds"]) / len(r["seconds"])
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * t, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": workload_config(shape, args.gpus),
        "overlaps_per_s": r["overlaps_per_s"],
        "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample", "host_cores")},
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


# ---------------------------------------------------------------------------------------------
def workload_shape(n_gpus: int, scale: float = 1.0):
    base = synth.CONFIGS[CONFIG_ID]
    f = n_gpus * scale
    return base if f == 1 else base.scaled(f, name="%sx%g" % (base.name, f))


def workload_config(shape, n_gpus):
    return {
        "workload": "BASELINE configs[1]: synthetic triploid 4.6 Mb genome, 30x per haplotype of "
                    "15 kb error-free reads, -l 1000, both strands%s"
                    % ("" if n_gpus == 1 else "; genome x%d (weak scaling, reads sharded by range)" % n_gpus),
        "shape": shape.name, "genome": shape.genome, "ploidy": shape.ploidy, "coverage": shape.cov,
        "read_len": shape.read_len, "min_length": shape.min_length,
        "reads": shape.n_reads, "entries": 2 * shape.n_reads, "forward_bases": shape.n_bases,
        "parallelism": "read-range x%d" % n_gpus,
        "l2": "inputs larger than L2 (ASCII reads %.0f MB per GPU vs 126 MB L2), no flush"
              % (shape.n_bases / n_gpus / 1e6),
    }


def run_b200(args):
    import torch
    import torch.distributed as dist
    from phasm_b200.overlapper import ExactOverlapper

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus %d needs torchrun with %d ranks (one process per GPU)"
                             % (args.gpus, args.gpus))
        args.gpus = world
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    shape = workload_shape(world, args.scale)
    l = shape.min_length
    R = shape.n_reads
    lo, hi = rank * R // world, (rank + 1) * R // world
    buf, off, names = synth.generate(shape, (lo, hi))
    n_bases_total = shape.n_bases

    ov = ExactOverlapper(local)
    for kv in args.opt:
        name, _, val = kv.partition("=")
        ov.set_option(name, int(val))
    rl = shape.read_len
    if lo:
        ov.add_remote(np.full(lo, rl, dtype=np.uint32))
    ov.add_reads_bulk(buf, off, names)
    if R - hi:
        ov.add_remote(np.full(R - hi, rl, dtype=np.uint32))
    if world > 1:
        uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            uid.copy_(torch.frombuffer(bytearray(ExactOverlapper.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(uid, 0)
        ov.comm_init(bytes(uid.cpu().numpy().tobytes()), rank, world)
    ov.upload()
    if world > 1:
        ov.comm_sync_ranges()

    def step_device():
        ov.overlaps_device(l)
        return ov.comm_gather_rows_device() if world > 1 else 0

    # ---- two untimed passes with the verifier's byte counters on (data-determined quantities):
    #      (1) every canonical candidate compared = the accounting SURVEY.md 8(d) defines;
    #      (2) with the strand-mirror halving the timed steps use = the bytes actually compared.
    user_mirror = 1
    for kv in args.opt:
        if kv.startswith("mirror="):
            user_mirror = int(kv.split("=")[1])
    ov.set_collect_stats(True)
    ov.set_option("mirror", 0)
    step_device()
    st_c = ov.stats()
    ov.set_option("mirror", user_mirror)
    step_device()
    st_x = ov.stats()
    ov.set_collect_stats(False)
    operand_bytes = st_c["operand_bytes"]
    candidates = st_c["candidates"]
    executed_bytes = st_x["operand_bytes"]

    # ---- device-resident: W warm-up + K timed steps ---------------------------------------
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()     # sampled from the warm-up on, so that a short timed region still sees samples
    for _ in range(args.warmup):
        step_device()
    barrier()
    stage = {k: 0.0 for k in ("ms_pack", "ms_anchor", "ms_scan", "ms_sort", "ms_verify", "ms_emit", "ms_device")}
    launches = lib_launches = 0
    ov.timer_start()
    for _ in range(args.steps):
        step_device()
        st = ov.stats()
        for k in stage:
            stage[k] += st[k]
        launches += st["launches"]
        lib_launches += st["lib_launches"]
    ms_total = ov.timer_stop()
    barrier()
    ms_total = max_over_ranks(ms_total)
    ms_step = ms_total / args.steps
    for k in stage:
        stage[k] /= args.steps
    stage_rank0 = dict(stage)
    if world > 1:   # slowest rank per stage (the step time is the max over ranks too)
        t = torch.tensor([stage[k] for k in sorted(stage)], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        stage = {k: float(v) for k, v in zip(sorted(stage), t.tolist())}
    st = ov.stats()
    n_rows = st["rows"] if rank == 0 else 0
    n_rows = int(sum_over_ranks(float(n_rows)))  # after the gather rank 0 holds all rows
    value = n_bases_total / (ms_step * 1e-3) / 1e9

    # ---- end to end with host buffers: H2D + device + D2H every step --------------------------
    def step_e2e():
        if world == 1:
            return len(ov.overlaps_arrays(l, copy=False))
        ov.upload()                      # ovl_upload always re-sends the host staging area
        ov.overlaps_device(l)
        return len(ov.comm_gather_rows(copy=False))

    for _ in range(min(args.warmup, 3)):
        step_e2e()
    e2e_steps = max(min(args.steps, 10), 1)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_e2e()
    torch.cuda.synchronize()
    t_e2e = max_over_ranks(time.perf_counter() - t0) / e2e_steps
    barrier()
    clocks = sampler.stop() if rank == 0 else None   # warm-up + timed region + e2e steps
    st_e = ov.stats()
    h2d = int(sum_over_ranks(float(st_e["ascii_bytes"] + 17 * st_e["entries"] + 4)))
    d2h = n_rows * 20
    e2e_value = n_bases_total / t_e2e / 1e9

    # ---- roofline of the dominant kernel ----------------------------------------------------
    peak, peak_src = measured_peak_gbs()
    n_local = int(off[-1])
    kern = {
        "k1_pack": (1.5 * n_local, stage["ms_pack"]),
        "k3_scan": (0.5 * n_local, stage["ms_scan"]),
        "k4_verify": (float(executed_bytes), stage["ms_verify"]),
    }
    dom = max(kern, key=lambda k: kern[k][1])
    per_kernel = {}
    for k, (b, ms) in kern.items():
        gbs = b / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
        per_kernel[k] = {"alg_bytes": b, "ms": ms, "GBps": gbs, "frac": gbs / peak}
    b_alg = (2.0 * n_local + 32.0 * st["anchors"] / world + operand_bytes
             + 32.0 * candidates + 20.0 * st_c["rows"])
    traffic, traffic_src = ncu_traffic()
    tmap = {"k1_pack": ("k1_pack",), "k3_scan": ("k3_probe",), "k4_verify": ("k4a_head", "k4b_tail")}
    for k in per_kernel:
        t = None
        if traffic:   # ncu names carry template arguments (k3_probe<6>): match on the prefix
            got = [sum(v for n, v in traffic.items() if n.startswith(p)) for p in tmap[k]]
            if all(x > 0 for x in got):
                t = sum(got)
        per_kernel[k]["traffic"] = t
        # measured DRAM bytes of the ncu capture / live kernel time: the fraction of the HBM peak
        # the kernel really draws (the algorithmic figure above counts operand bytes served by L1/L2)
        ms_k = per_kernel[k]["ms"]
        per_kernel[k]["dram_frac"] = (t / (ms_k * 1e-3) / 1e9 / peak) if (t and ms_k > 0) else None
    roofline = {
        "bound": "hbm", "kernel": dom, "achieved": per_kernel[dom]["GBps"], "peak": peak,
        "unit": "GB/s", "frac": per_kernel[dom]["frac"], "traffic": per_kernel[dom]["traffic"],
        "traffic_source": traffic_src, "dram_frac": per_kernel[dom]["dram_frac"],
        "peak_source": peak_src, "alg_bytes_per_launch": kern[dom][0],
        "ms_per_launch": kern[dom][1], "per_kernel": per_kernel,
        "k4_verify_canonical": {
            "note": "SURVEY 8(d) operand bytes of ALL canonical candidates / verify time; the timed "
                    "path compares only one candidate of each strand-mirror pair (alg_bytes above)",
            "alg_bytes": float(operand_bytes),
            "GBps": operand_bytes / (stage["ms_verify"] * 1e-3) / 1e9 if stage["ms_verify"] > 0 else 0.0,
            "candidates": candidates, "compared": st_x["candidates"]},
        "whole_path": {"alg_bytes": b_alg, "GBps": b_alg / (ms_step * 1e-3) / 1e9,
                       "frac": b_alg / (ms_step * 1e-3) / 1e9 / peak},
    }

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": workload_config(shape, world),
            "overlaps_per_s": n_rows / (ms_step * 1e-3), "rows": n_rows,
            "candidates": candidates, "verified": st["verified"], "hits": st["hits"],
            "mirror": st["mirror"], "anchor_stride": st["anchor_stride"], "stage_ms": stage, "stage_ms_rank0": stage_rank0,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "ms_per_step": 1e3 * t_e2e,
                    "overlaps_per_s": n_rows / t_e2e,
                    "ms_h2d": st_e["ms_h2d"], "ms_device": st_e["ms_device"], "ms_d2h": st_e["ms_d2h"],
                    "call": "ExactOverlapper.overlaps_arrays(l, copy=False) == ovl_overlaps: host ASCII in, rows out in the library's pinned host buffer"},
            "gpu_launches": launches, "lib_launches": lib_launches,
            "roofline": roofline, "clocks": clocks,
        }
        if world == 1 and not args.no_cpu:
            cb = cpu_reference_run(shape, args.cpu_sample, 1)
            line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample",
                                                       "host_cores", "overlaps_per_s")}
        print(json.dumps(line), flush=True)
    ov.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--scale", type=float, default=1.0, help="debug: scale the genome")
    ap.add_argument("--cpu-sample", type=int, default=4_000_000,
                    help="forward bases of the cpu_baseline sample (~15-20 s of reference CPU work)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--opt", action="append", default=[], metavar="NAME=VALUE",
                    help="debug: ovl_set_option tuning knob (e.g. anchor_stride=16, mirror=0)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    args.steps = max(args.steps, 1)
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
