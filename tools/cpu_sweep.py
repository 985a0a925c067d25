#!/usr/bin/env python
"""BASELINE.json config 1: the lab1 CPU build over a vector_size sweep (runs without a GPU),
plus the lab2 CPU filter on the two real test images.  Times the reference's own `main.c`
builds (baseline/_ref/bin, gcc as the reference README says) and this repository's CPU
binaries on the same box, same inputs, through the same stdin protocol; both print
`CPU execution time: <X ms>` around the compute loop only, which is the number compared
(median of `--reps`), and the whole-process wall time is given beside it.

    python tools/cpu_sweep.py --json profiles/cpu_sweep.json
"""
import argparse
import json
import re
import statistics
import subprocess
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
TIME_RE = re.compile(r"execution time: <([\d.]+) ms>")


def run(exe: Path, stdin_text: str):
    t0 = time.perf_counter()
    res = subprocess.run([str(exe)], input=stdin_text, capture_output=True, text=True, check=True)
    wall = (time.perf_counter() - t0) * 1e3
    m = TIME_RE.search(res.stdout[:200])
    return (float(m.group(1)) if m else None), wall, res.stdout


def lab1_input(n: int, rng) -> str:
    a = rng.uniform(-1e100, 1e100, n)
    b = rng.uniform(-1e100, 1e100, n)
    fmt = lambda v: " ".join(f"{x:.10e}" for x in v)  # noqa: E731
    return f"{n}\n{fmt(a)}\n{fmt(b)}\n"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--json", default="")
    ap.add_argument("--reps", type=int, default=5)
    args = ap.parse_args()
    ref = ROOT / "baseline" / "_ref" / "bin"
    rng = np.random.default_rng(5)
    rows = []
    for n in (1000, 10_000, 100_000, 1_000_000):
        text = lab1_input(n, rng)
        outs = {}
        for name, exe in (("reference main.c", ref / "lab1_cpu_exe"), ("ours", ROOT / "lab1/src/cpu_exe_to_plot")):
            if not exe.exists():
                continue
            ts, ws = [], []
            for _ in range(args.reps):
                t, w, out = run(exe, text)
                ts.append(t)
                ws.append(w)
            outs[name] = out.split("\n", 1)[1] if "\n" in out else out
            rows.append({"lab": "lab1", "config": f"n={n}", "impl": name, "kernel_ms_median": statistics.median(ts),
                         "process_wall_ms_median": round(statistics.median(ws), 2)})
        if len(outs) == 2:
            same = outs["reference main.c"].split() == outs["ours"].split()
            rows[-1]["output_identical_to_reference"] = same
    for img in ("lab2/test_data/lenna.data", "lab2/test_data/world_map.data"):
        src = ROOT / img
        if not src.exists():
            continue
        outs = {}
        for name, exe in (("reference main.c", ref / "lab2_cpu_exe"), ("ours", ROOT / "lab2/src/cpu_exe")):
            if not exe.exists():
                continue
            dst = Path("/tmp") / f"cpu_sweep_{name.split()[0]}_{src.name}"
            ts, ws = [], []
            for _ in range(args.reps):
                t, w, _ = run(exe, f"{src}\n{dst}\n")
                ts.append(t)
                ws.append(w)
            outs[name] = dst.read_bytes()
            rows.append({"lab": "lab2", "config": src.name, "impl": name, "kernel_ms_median": statistics.median(ts),
                         "process_wall_ms_median": round(statistics.median(ws), 2)})
        if len(outs) == 2:
            a = np.frombuffer(outs["reference main.c"], dtype=np.uint8)
            b = np.frombuffer(outs["ours"], dtype=np.uint8)
            rows[-1]["bytes_differing_from_reference_cpu"] = int((a != b).sum()) if a.size == b.size else -1
    for r in rows:
        print(json.dumps(r))
    if args.json:
        Path(args.json).write_text("\n".join(json.dumps(r) for r in rows) + "\n")


if __name__ == "__main__":
    sys.exit(main())
