#!/usr/bin/env python
"""Wall time of the lab2 contract binary, file to file, against the unmodified reference binary.

Both are driven through the same stdin protocol on the same /dev/shm files (fresh process per
run: process start + CUDA context + file read + H2D + kernel + D2H + file write), outputs are
compared byte for byte, and a bound is printed beside them: file size / page-cache copy bandwidth
(measured here with a plain read() into a reused buffer) + file size / PCIe bandwidth, each way.
Usage: python tools/binary_e2e.py [--size 16384] [--runs 3] [--json out.json]"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402


def wall(exe, stdin, runs, env):
    out = []
    for _ in range(runs):
        t0 = time.perf_counter()
        r = subprocess.run([str(exe)], input=stdin, capture_output=True, text=True, env=env)
        out.append((time.perf_counter() - t0) * 1e3)
        if r.returncode != 0:
            raise SystemExit(f"{exe}: {r.stderr[-400:]}")
    return out, float(r.stdout.split("<")[1].split(" ms>")[0])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=16384)
    ap.add_argument("--runs", type=int, default=3)
    ap.add_argument("--json", default=None)
    a = ap.parse_args()
    shm = Path("/dev/shm")
    src, out_ref, out_own = shm / "b200_e2e_in.data", shm / "b200_e2e_ref.data", shm / "b200_e2e_own.data"
    bench.write_synthetic_image(src, a.size, a.size, seed=3)
    nbytes = os.path.getsize(src)
    env = {**os.environ, "CUDA_VISIBLE_DEVICES": "0"}
    res = {"image": f"{a.size}x{a.size}", "file_bytes": nbytes}
    try:
        # page-cache copy bandwidth: read() the file into a reused 64 MiB buffer, single thread
        buf, t0 = bytearray(64 << 20), time.perf_counter()
        with open(src, "rb", buffering=0) as f:
            while f.readinto(buf):
                pass
        res["page_cache_read_GBps_1thread"] = round(nbytes / (time.perf_counter() - t0) / 1e9, 2)
        # process start + CUDA context alone: the smallest possible job
        tiny = shm / "b200_e2e_tiny.data"
        bench.write_synthetic_image(tiny, 4, 4, seed=1)
        t, _ = wall(ROOT / "lab2/src/to_plot_cuda_exe", f"-1\n-1\n-1\n-1\n{tiny}\n{shm / 'b200_e2e_tiny_out.data'}\n", 3, env)
        res["own_binary_tiny_image_wall_ms"] = round(statistics.median(t), 1)
        ref = ROOT / "baseline/_ref/bin/lab2_to_plot"
        if ref.exists():
            wall(ref, f"32\n32\n16\n16\n{src}\n{out_ref}\n", 1, env)
            t, k = wall(ref, f"32\n32\n16\n16\n{src}\n{out_ref}\n", a.runs, env)
            res["reference"] = {"wall_ms": [round(x, 1) for x in t], "median_ms": round(statistics.median(t), 1), "kernel_ms": k}
        wall(ROOT / "lab2/src/to_plot_cuda_exe", f"-1\n-1\n-1\n-1\n{src}\n{out_own}\n", 1, env)
        t, k = wall(ROOT / "lab2/src/to_plot_cuda_exe", f"-1\n-1\n-1\n-1\n{src}\n{out_own}\n", a.runs, env)
        res["own"] = {"wall_ms": [round(x, 1) for x in t], "median_ms": round(statistics.median(t), 1), "kernel_ms": k}
        if ref.exists():
            same = subprocess.run(["cmp", "-s", str(out_ref), str(out_own)]).returncode == 0
            res["outputs_byte_identical"] = same
            res["speedup_wall"] = round(res["reference"]["median_ms"] / res["own"]["median_ms"], 2)
        pcie = 50.0   # GB/s each way, measured by bench.py's e2e arm on this pool
        bound_ms = res["own_binary_tiny_image_wall_ms"] + 2 * nbytes / (res["page_cache_read_GBps_1thread"] * 1e6) \
            + 2 * nbytes / (pcie * 1e6)
        res["bound_ms_startup_plus_2x_pagecache_plus_2x_pcie"] = round(bound_ms, 1)
        res["own_over_bound"] = round(res["own"]["median_ms"] / bound_ms, 2)
    finally:
        for p in shm.glob("b200_e2e_*"):
            p.unlink()
    print(json.dumps(res, indent=1))
    if a.json:
        Path(a.json).write_text(json.dumps(res, indent=1))
    if res.get("outputs_byte_identical") is False:
        raise SystemExit("outputs differ")


if __name__ == "__main__":
    main()
