"""Host-side checks of the measurement plumbing (no GPU): the committed ncu capture that
bench.py's `roofline.traffic` comes from, and the reference arm's contract keys."""
import json
import os
import subprocess
import sys

import helpers

ROOT = helpers.ROOT


def test_committed_traffic_capture_covers_the_hot_kernels():
    sys.path.insert(0, ROOT)
    import bench
    import glob
    newest = sorted(glob.glob(os.path.join(ROOT, "profiles", "*_traffic.json")))[-1]
    with open(newest) as f:
        cap = json.load(f)
    n_cap, stride_cap = int(cap["n_gpus"]), int(cap["anchor_stride"])
    opts_cap = list(cap.get("options", []))
    traffic, src = bench.ncu_traffic(n_cap, stride_cap, opts_cap)
    assert traffic and src == os.path.basename(newest)
    # ... and of the same option set (a capture of another K1 grid is not this run's)
    t, why = bench.ncu_traffic(n_cap, stride_cap, opts_cap + ["pack_blocks_per_sm=3"])
    assert t is None and "options" in why
    for prefixes in (("k1_pack",), ("k3_filter", "k3_probe"), ("k4a_head",), ("k4b_tail",)):
        got = [v for n, v in traffic.items() if n.startswith(prefixes)]
        assert got and all(v > 1e6 for v in got), prefixes   # bytes per launch of the bench workload
    # the match rule: a capture is only pasted onto a run of the same GPU count
    # and anchor stride; anything else reports traffic = null and says why
    for n_gpus, stride in ((n_cap + 7, stride_cap), (n_cap, stride_cap * 2), (n_cap + 1, 16)):
        t, why = bench.ncu_traffic(n_gpus, stride, opts_cap)
        assert t is None and "capture" in why, (n_gpus, stride, why)
    # the launch list and the full summary of the same round are committed beside it
    tag = src.split("_traffic.json")[0]
    for suffix in ("_launches.csv", "_launches_summary.txt", "_ncu_full_summary.csv", "_bench.json"):
        assert os.path.exists(os.path.join(ROOT, "profiles", tag + suffix)), suffix
    with open(os.path.join(ROOT, "profiles", tag + "_bench.json")) as f:
        line = json.loads(f.read().strip().splitlines()[-1])
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better",
                "scaling", "vs_baseline", "dtype", "data", "config", "e2e", "gpu_launches", "roofline",
                "cpu_baseline", "clocks"):
        assert key in line, key
    assert line["roofline"]["traffic"] and 0 < line["roofline"]["dram_frac"] < 1
    assert line["gpu_launches"] > 0 and line["e2e"]["h2d_bytes_per_step"] > 0


def test_reference_arm_nonzero_rank_exits_quietly():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2",
                        "--steps", "1", "--warmup", "0"], capture_output=True, text=True, env=env, timeout=120)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_full_reference_digest_is_order_independent_and_matches_the_golden_helper():
    """bench.py's whole-row-set check against the recorded full-size run of the compiled reference
    (tests/golden/config<k>_full.json) uses the digest of tests/helpers.py::rows_digest."""
    import numpy as np
    sys.path.insert(0, ROOT)
    import bench
    rng = np.random.default_rng(3)
    rows = rng.integers(0, 50, size=(5000, 5), dtype=np.uint32)
    rows[100] = rows[7]                      # duplicate rows are part of the multiset
    want = helpers.rows_digest(rows)["sha256"]
    assert bench.full_rows_digest(rows) == want
    assert bench.full_rows_digest(rows[rng.permutation(len(rows))]) == want
    assert bench.full_rows_digest(rows[:0]) == helpers.rows_digest(rows[:0])["sha256"]
    for k in (1, 2):
        with open(os.path.join(ROOT, "tests", "golden", "config%d_full.json" % k)) as f:
            g = json.load(f)
        assert g["t2_equals_t0"] and g["t0"]["rows"] > 0 and len(g["t0"]["sha256"]) == 64
    rec = bench.recorded_full_reference()
    assert rec and rec["forward_bases"] == 414000000 and rec["cores_used"] == 1


def test_reference_arm_times_the_cpu_baseline_sample_whatever_the_step_count():
    """VERDICT r2 "next" 6c: `--impl reference` and the GPU arm's `cpu_baseline` leg time the SAME fixed
    genome-window sample (--cpu-sample, 4 Mbases), independent of --steps / --warmup, so one record never
    carries two different CPU numbers (and the GPU arm's `equal_work` key compares like with like)."""
    sys.path.insert(0, ROOT)
    import bench
    from phasm_b200 import synth
    shape = bench.workload_shape(1)
    _, off, names, _ = synth.window_subsample(shape, 4_000_000)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                        "--warmup", "0"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["config"]["reads"] == len(names)
    assert line["config"]["forward_bases"] == int(off[-1])
    assert line["cpu_baseline"]["sample"].startswith("%d reads / %d forward bases" % (len(names), int(off[-1])))
    assert line["e2e"]["value"] == line["value"] and line["e2e"]["h2d_bytes_per_step"] == 0
    src = open(os.path.join(ROOT, "bench.py")).read()
    assert '"equal_work"' in src and "args.steps + args.warmup" not in src
