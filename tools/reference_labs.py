#!/usr/bin/env python
"""Device ms of the UNMODIFIED reference lab3 binary (baseline/_ref/bin/lab3_to_plot) on
the BASELINE config: synthetic 16384 x 16384 RGBA image, 32 classes x 4096 sample points,
its own launch shape <<<256,256>>>, next to this build's binary on the same input.
Writes gpurun_out/reference_labs.json.  With --sweep: the reference binary over twelve launch shapes,
one run each, into gpurun_out/reference_lab3_sweep.json."""
import json
import os
import subprocess
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent


def main():
    size = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 16384
    shm = Path("/dev/shm")
    src, out = shm / "lab3_in.data", shm / "lab3_out.data"
    rng = np.random.default_rng(3)
    with open(src, "wb") as f:
        f.write(np.array([size, size], dtype=np.int32).tobytes())
        block = rng.integers(0, 256, size=1024 * size * 4, dtype=np.uint8).tobytes()
        for _ in range(size // 1024):
            f.write(block)
    spec = "32\n" + "\n".join(
        "4096 " + " ".join(f"{x} {y}" for x, y in zip(rng.integers(0, size, 4096), rng.integers(0, size, 4096)))
        for _ in range(32))
    res = {"config": f"{size}x{size}, 32 classes x 4096 points"}
    if "--sweep" in sys.argv:   # the reference binary over launch shapes <<<blocks, threads>>>, one run each
        ref = ROOT / "baseline/_ref/bin/lab3_to_plot"
        rows = []
        for blocks, threads in ((256, 256), (592, 256), (1184, 256), (2368, 256), (4736, 256), (16384, 256),
                                (65535, 256), (592, 512), (1184, 512), (592, 1024), (1184, 1024), (148, 1024)):
            r = subprocess.run([str(ref)], input=f"{blocks}\n{threads}\n{src}\n{out}\n{spec}", capture_output=True,
                               text=True)
            ms = float(r.stdout.split("<")[1].split(" ms>")[0]) if r.returncode == 0 and "<" in r.stdout else None
            rows.append({"blocks": blocks, "threads": threads, "kernel_ms": ms})
            print(rows[-1], flush=True)
        r = subprocess.run([str(ROOT / "lab3/src/to_plot_cuda_exe")], input=f"-1\n-1\n{src}\n{out}\n{spec}",
                           capture_output=True, text=True)
        res["reference_sweep"] = rows
        res["reference_best"] = min((x for x in rows if x["kernel_ms"]), key=lambda x: x["kernel_ms"])
        res["ours_tuned_ms"] = float(r.stdout.split("<")[1].split(" ms>")[0]) if r.returncode == 0 else None
        for p in (src, out):
            p.unlink(missing_ok=True)
        (ROOT / "gpurun_out").mkdir(exist_ok=True)
        (ROOT / "gpurun_out/reference_lab3_sweep.json").write_text(json.dumps(res, indent=1))
        print(json.dumps({k: res[k] for k in ("reference_best", "ours_tuned_ms")}))
        return
    for name, exe, dims in (("reference", ROOT / "baseline/_ref/bin/lab3_to_plot", "256\n256\n"),
                            ("ours_same_launch_shape", ROOT / "lab3/src/to_plot_cuda_exe", "256\n256\n"),
                            ("ours_tuned", ROOT / "lab3/src/to_plot_cuda_exe", "-1\n-1\n")):
        if not exe.exists():
            continue
        times = []
        for _ in range(3):
            t0 = time.time()
            r = subprocess.run([str(exe)], input=f"{dims}{src}\n{out}\n{spec}", capture_output=True, text=True)
            if r.returncode != 0:
                res[name] = {"error": r.stderr[-300:]}
                break
            times.append(float(r.stdout.split("<")[1].split(" ms>")[0]))
            wall = time.time() - t0
        else:
            res[name] = {"kernel_ms": times, "median_ms": sorted(times)[1], "wall_s_last": round(wall, 1)}
            res[name + "_sha"] = subprocess.run(["sha256sum", str(out)], capture_output=True, text=True).stdout.split()[0]
    for p in (src, out):
        p.unlink(missing_ok=True)
    (ROOT / "gpurun_out").mkdir(exist_ok=True)
    (ROOT / "gpurun_out/reference_labs.json").write_text(json.dumps(res, indent=1))
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
