#!/usr/bin/env python
"""BASELINE config 2 on a GPU box: the lab2 (and lab1/lab3) harness sweeps of the
reference's reports, run for this build's binaries and for the unmodified reference
binaries (baseline/_ref/bin, linked under <lab>/ref/ so that run_test.py finds the lab
name).  Writes profiles/harness_sweep.json with the median device ms per launch shape.
Usage: python tools/harness_sweep.py [--k_times 12]"""
import argparse
import json
import os
import subprocess
import sys
from pathlib import Path

import pandas as pd

ROOT = Path(__file__).resolve().parent.parent
LAB2_SIZES = "[[[16,16],[1024,1024]], [[16,16],[32,32]], [[2,2],[16,16]], [[32,32],[16,16]], [[32,32],[64,64]]]"
LAB1_SIZES = "[[1,32],[4,64],[32,128],[512,512],[1024,1024]]"
LAB3_SIZES = "[[256,256],[512,512],[1024,1024]]"


def link_reference(lab, name):
    src = ROOT / "baseline/_ref/bin" / name
    if not src.exists():
        return None
    d = ROOT / lab / "ref"
    d.mkdir(exist_ok=True)
    dst = d / f"ref_{name}"
    if dst.exists() or dst.is_symlink():
        dst.unlink()
    os.symlink(src, dst)
    return dst


def sweep(binary, sizes, k_times, extra=(), env=None):
    cmd = [sys.executable, "run_test.py", "--binary_path_cuda", str(binary), "--k_times", str(k_times),
           "--kernel_sizes", sizes, "--metadata_columns2plot", "[]", *extra]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=1800, env={**os.environ, **(env or {})})
    stem = Path(binary).stem
    csv = Path(binary).parent / f"stats_{stem}.csv"
    if r.returncode != 0 or "SUCCESS!" not in r.stdout or not csv.exists():
        return {"error": (r.stdout[-400:] + r.stderr[-400:])}
    df = pd.read_csv(csv)
    out = {}
    for ks, g in df.groupby("kernel_size"):
        out[ks] = {"median_ms": round(float(g["time_kernel_exe_ms"].median()), 5), "samples": int(g.shape[0]),
                   "verified": bool(g["test_verification_result"].all())}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--k_times", type=int, default=12)
    a = ap.parse_args()
    res = {}
    # (lab, our binary, reference binary, sizes, k_times, extra processor args)
    plans = [("lab2", "to_plot_cuda_exe", "lab2_to_plot", LAB2_SIZES, a.k_times, ()),
             ("lab3", "to_plot_cuda_exe", "lab3_to_plot", LAB3_SIZES, 4, ()),
             ("lab1", "cuda_exe_to_plot", "lab1_to_plot", LAB1_SIZES, 3,
              ("--min_vector_size", "10000", "--max_vector_size", "10001"))]
    out = ROOT / "gpurun_out" / "harness_sweep.json"
    out.parent.mkdir(exist_ok=True)
    for lab, ours, ref, sizes, k, extra in plans:
        # default = one cold launch, timed like the reference times itself; warm = after a warm-up
        res[f"{lab}.ours.cold"] = sweep(ROOT / lab / "src" / ours, sizes, k, extra)
        res[f"{lab}.ours.warm"] = sweep(ROOT / lab / "src" / ours, sizes, k, extra, {"B200LABS_WARM": "1"})
        link = link_reference(lab, ref)
        if link is not None:
            res[f"{lab}.reference"] = sweep(link, sizes, k, extra)
        out.write_text(json.dumps(res, indent=1))      # partial results survive a timeout
        print(lab, "done", flush=True)
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
