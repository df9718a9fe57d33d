#!/usr/bin/env python
"""bench.py -- throughput of the B200 overlapper on BASELINE.json's metric.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

Metric  : Gbases scanned/s (forward input bases / time) and overlaps/s, all-vs-all exact overlap.
Workload: BASELINE.json configs[1] -- synthetic triploid 4.6 Mb genome, 30x per haplotype of
          15 kb error-free reads, -l 1000 (414 Mbases, 27 600 reads, 55 200 entries) on ONE
          B200.  At N GPUs the genome is N times as long (weak scaling: every rank owns the
          same number of reads, sharded by read range).
Step    : one pass of the hot path (pack -> anchors -> scan -> verify -> sort -> emit) over the
          whole read set.  On N GPUs the rows end the step where they are emitted: sharded by A
          range over the ranks' HBM, like the reads (rank order = canonical order).  Pulling them
          into rank 0's HBM (what round 1 did inside the step) is ovl_fetch_rows on rank 0; its
          time is reported separately as `gather_to_rank0_ms`.
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
# the library keeps 9 streams busy per pass; the default of 8 hardware queues would alias two of them
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
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


def ncu_traffic(n_gpus: int, stride: int, options=()):
    """Per-kernel DRAM bytes per launch from the newest committed `ncu --set full` capture
    (profiles/<tag>_traffic.json, written by tools/profsum.py) -- but only if that capture was
    taken on THIS workload: same GPU count and anchor stride (a capture of the 1-GPU stride-64
    scan says nothing about a 4-GPU run).  Otherwise (None, why)."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "*_traffic.json")))
    if not files:
        return None, "no capture committed"
    try:
        with open(files[-1]) as f:
            d = json.load(f)
        if int(d.get("n_gpus", 1)) != n_gpus or int(d.get("anchor_stride", 64)) != stride:
            return None, "%s is a capture of n_gpus=%s stride=%s, this run is n_gpus=%d stride=%d" % (
                os.path.basename(files[-1]), d.get("n_gpus", 1), d.get("anchor_stride", 64), n_gpus, stride)
        # ... and with the same ovl_set_option set (launch shapes follow the options: VERDICT r2 weak 9(i))
        if sorted(d.get("options", [])) != sorted(options):
            return None, "%s is a capture with options %s, this run has %s" % (
                os.path.basename(files[-1]), sorted(d.get("options", [])), sorted(options))
        return {n: v["dram_bytes_per_launch"] for n, v in d["kernels"].items()}, os.path.basename(files[-1])
    except Exception as e:   # noqa: BLE001
        return None, "unreadable capture: %s" % e


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
        "overlaps_per_s": rows / t, "rows": rows, "seconds": secs, "bases": bases, "reads": n_reads,
        "_sample": (buf, off, names),   # not JSON: the reads themselves, for the equal-work GPU leg
    }


def recorded_full_reference():
    """The compiled reference's one recorded run on the FULL bench workload (build host, one core):
    tests/golden/config<k>_full.json, written by tests/golden/make_config1_full.py.  Not a live
    measurement of this run -- reported beside the live sample so that an equal-work figure exists."""
    p = os.path.join(ROOT, "tests", "golden", "config%d_full.json" % CONFIG_ID)
    try:
        with open(p) as f:
            g = json.load(f)
        return {"seconds": g["t0_overlaps_seconds"], "gbases_per_s": g["t0_gbases_per_s"],
                "overlaps_per_s": g["t0_overlaps_per_s"], "rows": g["t0"]["rows"], "peak_rss_gb": g["t0_peak_rss_gb"],
                "forward_bases": g["forward_bases"], "host_cores": g["host"]["cores"], "cores_used": 1,
                "measured_on": "build host (not this box), recorded once", "source": os.path.relpath(p, ROOT)}
    except Exception:   # noqa: BLE001
        return None


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    shape = workload_shape(args.gpus, args.scale)
    # ONE CPU sample for both CPU legs (VERDICT r2 "next" 6c): the same fixed genome window the
    # cpu_baseline leg of the GPU arm times (--cpu-sample, 4 Mbases = ~5.6 s of the reference per
    # call), whatever --steps / --warmup say; 20 + 3 calls end in about two minutes.
    sample = max(int(args.cpu_sample), 4 * shape.read_len)
    for _ in range(args.warmup):
        cpu_reference_run(shape, sample, 1)
    r = cpu_reference_run(shape, sample, max(args.steps, 1))
    t = sum(r["seconds"]) / len(r["seconds"])
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * t, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": dict(workload_config(shape, args.gpus),
                       sampled_from={"reads": shape.n_reads, "entries": 2 * shape.n_reads,
                                     "forward_bases": shape.n_bases},
                       reads=r["reads"], entries=2 * r["reads"], forward_bases=r["bases"],
                       note="the reference needs ~40 min and 28-43 GB per call on the full workload "
                            "(BASELINE.md section 2): it is timed on a genome-window sample of it, so "
                            "any ratio against this line compares per-base throughput across problem "
                            "sizes (cross-size; the reference is super-linear, so it is under-stated)"),
        "overlaps_per_s": r["overlaps_per_s"],
        "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample", "host_cores")},
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "reference_full_workload_recorded": recorded_full_reference() if args.gpus == 1 else None,
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
        "output": "rows in HBM" if n_gpus == 1 else
                  "rows in HBM, sharded by A range over the ranks like the reads (rank order = canonical order)",
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
        # one pass; on N > 1 ranks every rank ends with the rows of its A range in its HBM
        return ov.overlaps_device(l)

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
    rows_canonical = int(sum_over_ranks(float(st_c["rows_local"])))

    # ---- device-resident: W warm-up + K timed steps ---------------------------------------
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()     # sampled from the warm-up on, so that a short timed region still sees samples
    for _ in range(args.warmup):
        step_device()
    barrier()
    stage_keys = ("ms_pack", "ms_anchor", "ms_scan", "ms_sort", "ms_verify", "ms_emit", "ms_gather", "ms_device")
    stage = {k: 0.0 for k in stage_keys}
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
    n_rows = int(sum_over_ranks(float(st["rows_local"])))
    value = n_bases_total / (ms_step * 1e-3) / 1e9

    # ---- rows of all ranks into rank 0's HBM (device pull over NVLink), timed on its own ----------
    gather_ms = None
    if world > 1:
        if rank == 0:
            ov.comm_gather_rows_device()
            gather_ms = ov.stats()["ms_gather"]
        barrier()

    # ---- parity, outside the timed region (rank 0 checks the rows of all ranks) -------------------
    parity = check_parity(ov, shape, world, rank, n_rows, args.parity_bases) if not args.no_parity else None
    barrier()

    # ---- end to end with host buffers: H2D + device + D2H every step --------------------------
    # 1 GPU: ovl_overlaps (pinned host ASCII in, chunked and packed under the copy; rows out).
    # N GPUs: ovl_overlaps_local on every rank (the same, and every rank brings its own rows to its
    # host over its own PCIe link).
    def step_e2e():
        if world == 1:
            return len(ov.overlaps_arrays(l, copy=False))
        return len(ov.overlaps_local_arrays(l, copy=False))

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
    h2d = int(sum_over_ranks(float(st_e["ascii_bytes"])))
    d2h = n_rows * 20
    e2e_value = n_bases_total / t_e2e / 1e9
    e2e_parts = {k: max_over_ranks(float(st_e[k])) for k in ("ms_h2d", "ms_exposed", "ms_d2h", "ms_device")}

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
    # whole path: per rank.  "executed" counts the operand bytes the verifier really
    # compared (one candidate of each strand-mirror pair); "canonical" all canonical candidates.
    fixed = 2.0 * n_local + 32.0 * st["anchors"] / world + 20.0 * rows_canonical / world
    b_exec = fixed + executed_bytes + 32.0 * st_x["candidates"]
    b_canon = fixed + operand_bytes + 32.0 * candidates
    traffic, traffic_src = ncu_traffic(world, int(st["anchor_stride"]), args.opt)
    tmap = {"k1_pack": ("k1_pack",), "k3_scan": ("k3_filter", "k3_resolve"), "k4_verify": ("k4a_head", "k4b_tail")}
    for k in per_kernel:
        t = None
        if traffic:   # ncu names carry template arguments (k3_filter<6>): match on the prefix
            got = [sum(v for n, v in traffic.items() if n.startswith(p)) for p in tmap[k]]
            if all(x > 0 for x in got):
                t = sum(got)
        per_kernel[k]["traffic"] = t
        # measured DRAM bytes of the ncu capture / live kernel time: the fraction of the HBM peak
        # the kernel really draws (the algorithmic figure above counts operand bytes served by L1/L2)
        ms_k = per_kernel[k]["ms"]
        per_kernel[k]["dram_frac"] = (t / (ms_k * 1e-3) / 1e9 / peak) if (t and ms_k > 0) else None
    # The fraction of the HBM roofline is a statement about DRAM traffic: when a matching ncu capture
    # exists it is (captured DRAM bytes per launch / live kernel time) / peak.  The algorithmic figure
    # (operand bytes / time) can exceed the peak because part of the operands is served by L1/L2; it is
    # reported beside it, labelled, and never as a fraction above 1.
    alg_frac = per_kernel[dom]["frac"]
    dram_frac = per_kernel[dom]["dram_frac"]
    if dram_frac is not None:
        frac, basis = dram_frac, "dram: ncu dram bytes per launch (%s) / live launch time / peak" % traffic_src
    else:
        frac = min(alg_frac, 1.0)
        basis = "algorithmic bytes / live launch time / peak" + (
            ", clamped at 1 (operands partly L1/L2-served; no matching ncu capture)" if alg_frac > 1.0 else "")
    roofline = {
        "bound": "hbm", "kernel": dom, "achieved": per_kernel[dom]["GBps"], "peak": peak,
        "unit": "GB/s", "frac": frac, "frac_basis": basis, "frac_algorithmic": alg_frac,
        "achieved_note": "algorithmic operand bytes / time (SURVEY 8(d)); above the HBM peak means L1/L2-served operands, "
                         "not a faster-than-peak kernel",
        "achieved_dram": (per_kernel[dom]["traffic"] / (kern[dom][1] * 1e-3) / 1e9) if (per_kernel[dom]["traffic"] and kern[dom][1] > 0) else None,
        "traffic": per_kernel[dom]["traffic"],
        "traffic_source": traffic_src, "dram_frac": per_kernel[dom]["dram_frac"],
        "peak_source": peak_src, "alg_bytes_per_launch": kern[dom][0],
        "ms_per_launch": kern[dom][1], "per_kernel": per_kernel,
        "k4_verify_canonical": {
            "note": "SURVEY 8(d) operand bytes of ALL canonical candidates / verify time; the timed "
                    "path compares only one candidate of each strand-mirror pair (alg_bytes above)",
            "alg_bytes": float(operand_bytes),
            "GBps": operand_bytes / (stage["ms_verify"] * 1e-3) / 1e9 if stage["ms_verify"] > 0 else 0.0,
            "candidates": candidates, "compared": st_x["candidates"]},
        "whole_path": {"alg_bytes": b_exec, "GBps": b_exec / (ms_step * 1e-3) / 1e9,
                       "frac": b_exec / (ms_step * 1e-3) / 1e9 / peak,
                       "basis": "executed bytes (rank 0's share; one candidate of each strand-mirror pair compared)",
                       "canonical": {"alg_bytes": b_canon, "frac": b_canon / (ms_step * 1e-3) / 1e9 / peak}},
    }

    if rank == 0:
        ms_e2e = 1e3 * t_e2e
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": workload_config(shape, world), "options": sorted(args.opt),
            "overlaps_per_s": n_rows / (ms_step * 1e-3), "rows": n_rows,
            "candidates": candidates, "verified": st["verified"], "hits": st["hits"],
            "scan_survivors": st["scan_survivors"], "big_buckets": st["big_buckets"], "retries": st["retries"],
            "mirror": st["mirror"], "anchor_stride": st["anchor_stride"], "stage_ms": stage, "stage_ms_rank0": stage_rank0,
            "parity": parity, "gather_to_rank0_ms": gather_ms,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e,
                    "overlaps_per_s": n_rows / t_e2e,
                    "ms_h2d": e2e_parts["ms_h2d"], "ms_device_after_upload": e2e_parts["ms_exposed"],
                    "ms_d2h": e2e_parts["ms_d2h"], "ms_device_incl_upload_wait": e2e_parts["ms_device"],
                    # (parts: last step, each the max over the ranks -- on N > 1 they come from different
                    #  ranks and need not add up to the wall time per step)
                    "ms_host_other": max(0.0, ms_e2e - e2e_parts["ms_h2d"] - e2e_parts["ms_exposed"] - e2e_parts["ms_d2h"]),
                    "upload": "chunked: pinned host ASCII -> HBM in pieces, every piece packed (K1) and "
                              "prefix-tabled under the next piece's copy",
                    "call": ("ExactOverlapper.overlaps_arrays(l, copy=False) == ovl_overlaps: host ASCII in, rows "
                             "out in the library's pinned host buffer") if world == 1 else
                            ("ExactOverlapper.overlaps_local_arrays(l, copy=False) == ovl_overlaps_local on every "
                             "rank: host ASCII in, the pass, every rank's own rows out over its own PCIe link "
                             "(concatenated in rank order they are the canonical row list)")},
            "gpu_launches": launches, "lib_launches": lib_launches,
            "roofline": roofline, "clocks": clocks,
        }
        if world == 1 and not args.no_cpu:
            cb = cpu_reference_run(shape, args.cpu_sample, 1)
            line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample",
                                                       "host_cores", "overlaps_per_s")}
            # beside the live sample: the reference's one recorded run on this very workload in full
            line["cpu_baseline"]["full_workload_recorded"] = recorded_full_reference()
            # ... and the GPU arm on that SAME sample, end to end with host buffers (ovl_overlaps):
            # the one same-config, same-box GPU / CPU figure of this line
            sbuf, soff, snames = cb["_sample"]
            ov2 = ExactOverlapper(local)
            ov2.add_reads_bulk(sbuf, soff, snames)
            for _ in range(3):
                rows2 = ov2.overlaps_arrays(l, copy=False)
            t_g = []
            for _ in range(10):
                t0 = time.perf_counter()
                rows2 = ov2.overlaps_arrays(l, copy=False)
                t_g.append(time.perf_counter() - t0)
            gpu_ms = 1e3 * sum(t_g) / len(t_g)
            cpu_ms = 1e3 * sum(cb["seconds"]) / len(cb["seconds"])
            line["equal_work"] = {
                "sample": cb["sample"], "reads": cb["reads"], "forward_bases": cb["bases"],
                "cpu_ms": cpu_ms, "cpu_kind": cb["kind"], "cpu_cores": cb["cores"], "cpu_rows": cb["rows"],
                "gpu_ms": gpu_ms, "gpu_ms_device": float(ov2.stats()["ms_device"]), "gpu_rows": int(len(rows2)),
                "gpu_call": "ovl_overlaps on the same reads: host ASCII in, rows out (wall clock, mean of 10 after 3 warm-up calls)",
                "rows_equal": int(len(rows2)) == int(cb["rows"]), "ratio": cpu_ms / gpu_ms if gpu_ms > 0 else None,
                "note": "a 4 Mbase job is launch- and latency-bound on the GPU (a few hundred reads on 148 SMs); "
                        "it is the equal-work figure, not the throughput figure",
            }
            ov2.close()
        print(json.dumps(line), flush=True)
    ok = 1 if (parity is None or parity.get("bit_exact", True)) else 0
    ov.close()
    if world > 1:
        t = torch.tensor([ok], device="cuda")
        dist.broadcast(t, 0)
        ok = int(t.item())
        dist.barrier()
        dist.destroy_process_group()
    return 0 if ok else 3


def full_rows_digest(rows) -> str:
    """sha256 of the rows sorted by (a, b, astart, aend, bend), as little-endian uint32 -- the digest
    tests/golden/make_config1_full.py records for the compiled reference (tests/helpers.py rows_digest)."""
    import hashlib
    arr = np.ascontiguousarray(rows, dtype=np.uint32).reshape(-1, 5)
    if len(arr):
        arr = np.ascontiguousarray(arr[np.lexsort((arr[:, 4], arr[:, 3], arr[:, 2], arr[:, 1], arr[:, 0]))])
    return hashlib.sha256(arr.tobytes()).hexdigest()


def check_parity(ov, shape, world, rank, n_rows, window_bases):
    """Driver-visible parity at every N, outside the timed region.  Rank 0
    fetches the gathered rows and checks (a) canonical order and an order-independent checksum,
    (b) the RESTRICTION property: the rows restricted to all reads that start in
    one genome window equal, bit for bit, the CPU oracle's rows (oracle/overlap_oracle.c, pinned on
    the compiled reference) on exactly those reads, (c) the strand-mirror invariant on all rows.
    The oracle is the checker here, never the thing measured."""
    if rank != 0:
        return None
    from oracle import port
    from oracle.bruteforce import both_strand_entries
    t0 = time.perf_counter()
    rows = ov.fetch_rows(copy=False)
    res = {"checked": True, "rows": int(len(rows)), "rows_expected_from_ranks": int(n_rows)}
    ok = len(rows) == n_rows
    r64 = rows.astype(np.uint64)
    m2 = np.uint64(2 * shape.n_reads)
    pair = r64[:, 0] * m2 + r64[:, 1]
    mix = (pair * np.uint64(0x9E3779B97F4A7C15)) ^ (r64[:, 2] << np.uint64(40)) ^ (r64[:, 3] << np.uint64(20)) ^ r64[:, 4]
    mix = mix * np.uint64(0xD6E8FEB86659FD93)
    res["checksum64"] = "%016x" % int(np.bitwise_xor.reduce(mix) ^ np.uint64(np.sum(mix, dtype=np.uint64))) if len(rows) else "0" * 16
    if len(rows) > 1:
        d_pair = pair[1:] >= pair[:-1]
        same = pair[1:] == pair[:-1]
        in_order = bool(np.all(d_pair) and np.all(r64[1:, 2][same] >= r64[:-1, 2][same]))
    else:
        in_order = True
    res["canonical_order"] = in_order
    ok = ok and in_order
    # (b) restriction
    wb, wo, wn, widx = synth.window_subsample(shape, int(window_bases))
    reads = [(n, wb[int(wo[i]):int(wo[i + 1])].tobytes().decode()) for i, n in enumerate(wn)]
    ebuf, eoff = port.entries_to_buffer(both_strand_entries(reads))
    want, _ = port.oracle_overlaps(ebuf, eoff, shape.min_length)
    remap = -np.ones(2 * shape.n_reads, dtype=np.int64)
    remap[2 * widx] = 2 * np.arange(len(widx))
    remap[2 * widx + 1] = 2 * np.arange(len(widx)) + 1
    keep = (remap[rows[:, 0]] >= 0) & (remap[rows[:, 1]] >= 0) if len(rows) else np.zeros(0, bool)
    f = np.array(rows[keep], dtype=np.int64)
    if len(f):
        f[:, 0] = remap[f[:, 0]]
        f[:, 1] = remap[f[:, 1]]
        f = f[np.lexsort((f[:, 4], f[:, 3], f[:, 2], f[:, 1], f[:, 0]))]
    exact = f.shape == want.shape and bool(np.array_equal(f.astype(np.uint32), want))
    res.update({"window_reads": int(len(widx)), "window_bases": int(wo[-1]), "window_rows": int(len(want)),
                "window_rows_gpu": int(len(f)), "bit_exact": bool(exact and ok),
                "oracle": "oracle/overlap_oracle.c (C restatement pinned on the compiled reference)"})
    # (c) strand-mirror invariant over all suffix-prefix rows (fixed read length)
    rl = shape.read_len
    sp = r64[r64[:, 3] == np.uint64(rl)] if len(rows) else r64
    if len(sp):
        k_fwd = np.sort((sp[:, 0] * m2 + sp[:, 1]) * np.uint64(rl + 1) + sp[:, 4])
        k_twin = np.sort(((sp[:, 1] ^ np.uint64(1)) * m2 + (sp[:, 0] ^ np.uint64(1))) * np.uint64(rl + 1) + sp[:, 4])
        res["mirror_invariant"] = bool(np.array_equal(k_fwd, k_twin))
        res["bit_exact"] = bool(res["bit_exact"] and res["mirror_invariant"])
    # (d) the whole row set against the UNMODIFIED reference at full size, when its digest is committed
    # for exactly this workload (tests/golden/config<k>_full.json: the compiled reference run once on
    # the build host, tests/golden/make_config1_full.py) -- N = 1 only: at N GPUs the genome is N times as long
    gold_path = os.path.join(ROOT, "tests", "golden", "config%d_full.json" % CONFIG_ID)
    if world == 1 and os.path.exists(gold_path):
        with open(gold_path) as f:
            gold = json.load(f)
        if gold.get("shape") == shape.name and gold.get("forward_bases") == shape.n_bases and "t0" in gold:
            digest = full_rows_digest(rows)
            res["reference_full"] = {"rows": gold["t0"]["rows"], "sha256_equal": digest == gold["t0"]["sha256"],
                                     "reference_seconds": gold.get("t0_overlaps_seconds"),
                                     "source": "tests/golden/config%d_full.json (oracle/_ref on the full workload)" % CONFIG_ID}
            res["bit_exact"] = bool(res["bit_exact"] and res["reference_full"]["sha256_equal"])
    res["seconds"] = time.perf_counter() - t0
    return res


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
    ap.add_argument("--no-parity", action="store_true", help="skip the parity check after the timed region")
    ap.add_argument("--parity-bases", type=float, default=40e6,
                    help="forward bases of the genome window the parity check runs the CPU oracle on")
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
