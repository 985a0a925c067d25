#!/usr/bin/env python
"""The UNMODIFIED reference lab2 binary (baseline/_ref/bin/lab2_to_plot) over its own tuning
knob - the launch geometry it reads from stdin - on large synthetic images, so that "vs the
reference" can be quoted against its BEST shape and not only its stock block 32x32 / grid 16x16.
Kernel ms are the binary's own cudaEvent line (median of --runs fresh processes per shape).
Usage: python tools/reference_sweep.py [--json gpurun_out/reference_sweep.json]"""
import argparse
import json
import os
import statistics
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402

BLOCKS = [(32, 8), (32, 16), (32, 32), (16, 16)]
GRIDS = [(16, 16), (64, 64), (256, 256), (1024, 1024)]


def kernel_ms(exe, stdin, runs, env):
    t = []
    for _ in range(runs):
        r = subprocess.run([str(exe)], input=stdin, capture_output=True, text=True, env=env)
        if r.returncode != 0:
            return None
        t.append(float(r.stdout.split("<")[1].split(" ms>")[0]))
    return statistics.median(t)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--runs", type=int, default=3)
    ap.add_argument("--json", default=str(ROOT / "gpurun_out" / "reference_sweep.json"))
    a = ap.parse_args()
    exe = ROOT / "baseline/_ref/bin/lab2_to_plot"
    own = ROOT / "lab2/src/to_plot_cuda_exe"
    shm = Path("/dev/shm")
    src, out = shm / "b200_sweep_in.data", shm / "b200_sweep_out.data"
    env = {**os.environ, "CUDA_VISIBLE_DEVICES": "0"}
    res = {}
    try:
        for label, w, h in (("16384x16384", 16384, 16384), ("65536x16384 (one 2^30-px chunk of 65536^2)", 65536, 16384)):
            bench.write_synthetic_image(src, w, h, seed=5)
            rows = []
            for bx, by in BLOCKS:
                for gx, gy in GRIDS:
                    ms = kernel_ms(exe, f"{bx}\n{by}\n{gx}\n{gy}\n{src}\n{out}\n", a.runs, env)
                    rows.append({"block": [bx, by], "grid": [gx, gy], "kernel_ms": ms,
                                 "GBps": round(8 * w * h / ms / 1e6, 1) if ms else None})
                    print(label, rows[-1], flush=True)
            ok = [r for r in rows if r["kernel_ms"]]
            best = min(ok, key=lambda r: r["kernel_ms"])
            stock = next(r for r in rows if r["block"] == [32, 32] and r["grid"] == [16, 16])
            ours = kernel_ms(own, f"-1\n-1\n-1\n-1\n{src}\n{out}\n", a.runs, env)
            res[label] = {"bytes_moved": 8 * w * h, "reference_stock_shape": stock, "reference_best_shape": best,
                          "this_build_cold_ms": ours, "this_build_GBps": round(8 * w * h / ours / 1e6, 1),
                          "speedup_vs_stock": round(stock["kernel_ms"] / ours, 2),
                          "speedup_vs_best": round(best["kernel_ms"] / ours, 2), "all_shapes": rows}
            Path(a.json).parent.mkdir(exist_ok=True)
            Path(a.json).write_text(json.dumps(res, indent=1))
    finally:
        for p in (src, out):
            if p.exists():
                p.unlink()
    print(json.dumps({k: {kk: vv for kk, vv in v.items() if kk != "all_shapes"} for k, v in res.items()}, indent=1))


if __name__ == "__main__":
    main()
