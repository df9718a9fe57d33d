#!/usr/bin/env python
"""Turn one GPU visit's scratch output (gpurun_out/<tag>/, written by tools/gpu_round.sh) into the
committed files under profiles/:

    <tag>_bench.json               the bench line of that visit
    <tag>_pytest_gpu.log           pytest -m gpu tail
    <tag>_nvidia_smi.txt
    <tag>_launches.csv             ncu --metrics gpu__time_duration.sum launch list
    <tag>_launches_summary.txt     per-kernel totals and shares of that list
    <tag>_ncu_full_summary.csv     one row per captured kernel of the `ncu --set full` report
    <tag>_traffic.json             per-kernel DRAM bytes per launch (bench.py's roofline.traffic)

Usage: python tools/profsum.py <tag>
"""
import collections
import csv
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

FULL_METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum",
    "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
    "launch__grid_size", "launch__block_size", "smsp__inst_executed.sum",
    "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
]

SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12,
         "ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}


def short(name: str) -> str:
    name = name.replace("ovl::", "")
    return name.split("(")[0].replace("void ", "")[:60]


def launches_summary(src: str, dst: str, cmdline: str) -> None:
    rows = []
    with open(src) as f:
        lines = [l for l in f if l.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r["Metric Unit"], 1.0)
        rows.append((r["Kernel Name"], v))
    tot = sum(v for _, v in rows) or 1.0
    agg = collections.OrderedDict()
    for k, v in rows:
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += v
    with open(dst, "w") as f:
        f.write("# %s\n# (cold-cache, serialised launches: compare SHARES, not absolutes)\n" % cmdline)
        f.write("%d launches, total us %.1f\n" % (len(rows), tot))
        for k, (n, v) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write("%-72s %5d %10.1f us avg %8.1f %5.1f%%\n" % (k[:72], n, v, v / n, 100 * v / tot))


def full_summary(rep: str, dst_csv: str, dst_json: str) -> None:
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, body = rows[0], rows[1], rows[2:]
    cols = [m for m in FULL_METRICS if m in hdr]
    ki = hdr.index("Kernel Name")
    per = collections.OrderedDict()
    with open(dst_csv, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["kernel"] + ["%s [%s]" % (m, units[hdr.index(m)]) for m in cols])
        for r in body:
            w.writerow([short(r[ki])] + [r[hdr.index(m)] for m in cols])
            def val(m):
                i = hdr.index(m)
                return float(r[i].replace(",", "")) * SCALE.get(units[i], 1.0)
            a = per.setdefault(short(r[ki]), {"launches": 0, "dram_bytes": 0.0, "us": 0.0})
            a["launches"] += 1
            a["dram_bytes"] += val("dram__bytes_read.sum") + val("dram__bytes_write.sum")
            a["us"] += val("gpu__time_duration.sum")
    res = {k: {"launches": a["launches"], "dram_bytes_per_launch": a["dram_bytes"] / a["launches"],
               "us_per_launch_under_ncu": a["us"] / a["launches"]} for k, a in per.items()}
    meta = {"n_gpus": 1, "anchor_stride": 64, "options": []}
    bj = os.path.join(os.path.dirname(rep), "bench.json")
    try:   # the bench line of the same visit says which workload the capture saw
        with open(bj) as f:
            line = json.loads([l for l in f if l.startswith("{")][-1])
        meta = {"n_gpus": int(line["n_gpus"]), "anchor_stride": int(line["anchor_stride"]),
                "options": list(line.get("options", []))}
    except Exception:
        pass
    with open(dst_json, "w") as f:
        json.dump({"n_gpus": meta["n_gpus"], "anchor_stride": meta["anchor_stride"], "options": meta.get("options", []), "source": "ncu --set full --clock-control none, python bench.py --steps 2 --warmup 3 --no-cpu "
                             "(BASELINE configs[1]); dram__bytes_read.sum + dram__bytes_write.sum per launch",
                   "kernels": res}, f, indent=1)
        f.write("\n")


def main():
    tag = sys.argv[1]
    src = os.path.join(ROOT, "gpurun_out", tag)
    dst = os.path.join(ROOT, "profiles")
    for name in ("bench.json", "pytest_gpu.log", "nvidia_smi.txt", "launches.csv"):
        p = os.path.join(src, name)
        if os.path.exists(p):
            shutil.copy(p, os.path.join(dst, "%s_%s" % (tag, name)))
    lc = os.path.join(src, "launches.csv")
    if os.path.exists(lc):
        launches_summary(lc, os.path.join(dst, tag + "_launches_summary.txt"),
                         "%s: ncu --metrics gpu__time_duration.sum --clock-control none -c 600, "
                         "python bench.py --steps 2 --warmup 3 --no-cpu" % tag)
    rep = os.path.join(src, "full.ncu-rep")
    if os.path.exists(rep):
        full_summary(rep, os.path.join(dst, tag + "_ncu_full_summary.csv"), os.path.join(dst, tag + "_traffic.json"))
    print("profiles/%s_* written" % tag)


if __name__ == "__main__":
    main()
