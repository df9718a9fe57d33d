#!/usr/bin/env python
"""T_e2e of the CLI ("FASTA file -> GFA2 file", and host materialisation separately).

Times, on one GPU and BASELINE config 2 (or --cfg / --scale):

  * `python -m phasm overlap -l L -o out.gfa reads.fasta`       native route (ovl_add_fasta + ovl_write_gfa)
  * `python -m phasm overlap --python-io ...`                    the reference's loop shape
    (phasm/cli/assembler.py:29-50 of the reference: FastaReader, add_sequence x2, overlaps(), gfa_line)
  * in process: ExactOverlapper.overlaps(l) -> list[tuple] (what src/phasm.cpp:2,15 converts to) against
    overlaps_arrays(l) -> (n, 5) array, both through the drop-in extension module

FASTA and GFA2 live on tmpfs (/dev/shm) so that no disk is measured.  Each subprocess time includes the
interpreter start, the CUDA context and the library load; `--stats json` gives the device / copy times
inside.  Prints one JSON object.  Measurement tool: the numbers go to profiles/, nothing here is a test.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from phasm_b200 import synth  # noqa: E402


def run_cli(extra, fasta, out, l):
    cmd = [sys.executable, "-m", "phasm", "overlap", "-l", str(l), "--stats", "json", "-o", out] + extra + [fasta]
    t0 = time.perf_counter()
    r = subprocess.run(cmd, capture_output=True, text=True, cwd=ROOT, timeout=1800)
    t = time.perf_counter() - t0
    if r.returncode != 0:
        raise SystemExit("phasm overlap failed: " + r.stderr[-2000:])
    st = {}
    for line in r.stderr.splitlines():
        if line.startswith('{"phasm_overlap_stats"'):
            st = json.loads(line)["phasm_overlap_stats"]
    return t, st


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cfg", type=int, default=2)
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--tmp", default="/dev/shm" if os.path.isdir("/dev/shm") else None)
    ap.add_argument("--skip-python-io", action="store_true")
    args = ap.parse_args()
    sh = synth.CONFIGS[args.cfg]
    if args.scale != 1.0:
        sh = sh.scaled(args.scale)
    buf, off, names = synth.generate(sh)
    res = {"shape": sh.name, "reads": len(names), "forward_bases": int(off[-1]), "min_length": sh.min_length}
    with tempfile.TemporaryDirectory(dir=args.tmp) as d:
        fasta = os.path.join(d, "reads.fasta")
        t0 = time.perf_counter()
        with open(fasta, "wb") as f:
            for i, n in enumerate(names):
                f.write(b">" + n.encode() + b"\n")
                f.write(buf[int(off[i]):int(off[i + 1])].tobytes())
                f.write(b"\n")
        res["fasta_bytes"] = os.path.getsize(fasta)
        res["s_write_fasta"] = time.perf_counter() - t0
        out_n, out_p = os.path.join(d, "native.gfa"), os.path.join(d, "python.gfa")
        ts = []
        for _ in range(3):
            t, st = run_cli([], fasta, out_n, sh.min_length)
            ts.append(t)
        res["native"] = {"s_wall_runs": ts, "s_wall": min(ts), "gfa_bytes": os.path.getsize(out_n),
                         "gbases_per_s": int(off[-1]) / min(ts) / 1e9,
                         "stats": {k: st.get(k) for k in ("rows", "ms_h2d", "ms_device", "ms_d2h", "ms_total", "launches")}}
        if not args.skip_python_io:
            t, st = run_cli(["--python-io"], fasta, out_p, sh.min_length)
            res["python_io"] = {"s_wall": t, "gfa_bytes": os.path.getsize(out_p), "gbases_per_s": int(off[-1]) / t / 1e9,
                                "stats": {k: st.get(k) for k in ("rows", "ms_h2d", "ms_device", "ms_d2h", "ms_total")}}
            with open(out_n, "rb") as f:
                a = f.read()
            with open(out_p, "rb") as f:
                b = f.read()
            res["routes_write_the_same_bytes"] = a == b
            if a != b:   # same lines, another E-line order?
                res["routes_write_the_same_lines"] = sorted(a.splitlines()) == sorted(b.splitlines())
        # interpreter + CUDA context + library load alone (an empty FASTA)
        empty = os.path.join(d, "empty.fasta")
        open(empty, "w").close()
        t, _ = run_cli([], empty, os.path.join(d, "empty.gfa"), sh.min_length)
        res["s_wall_empty_input"] = t

    # ---- host materialisation through the drop-in module (a9) -------------------------------------
    import __graft_entry__ as g
    g.build_library()
    g.build_pybind()
    import phasm.overlapper as mod
    ov = mod.ExactOverlapper()
    for i, n in enumerate(names):
        ov.add_read(n, buf[int(off[i]):int(off[i + 1])].tobytes().decode())
    ov.overlaps_arrays(sh.min_length)
    t0 = time.perf_counter()
    arr = ov.overlaps_arrays(sh.min_length)
    t_arr = time.perf_counter() - t0
    t0 = time.perf_counter()
    lst = ov.overlaps(sh.min_length)
    t_lst = time.perf_counter() - t0
    res["materialise"] = {"rows": int(len(arr)), "s_overlaps_arrays": t_arr, "s_overlaps_list_of_tuples": t_lst,
                          "s_tuples_only": t_lst - t_arr, "us_per_row_tuple": 1e6 * (t_lst - t_arr) / max(len(lst), 1),
                          "same_count": len(lst) == len(arr)}
    print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
