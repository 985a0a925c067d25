#!/usr/bin/env python
"""Single-GPU kernel micro-benchmarks (device ms by CUDA events, effective GB/s
against MEASURED_PEAKS.json).  Usage: python tools/microbench.py [--quick] [--json out.json]"""
import argparse
import json
import statistics
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from cuda_mpi_openmp_b200 import _native as N, ops  # noqa: E402
from cuda_mpi_openmp_b200.utils import cuda_time_ms  # noqa: E402


def peak_gbs():
    try:
        return json.loads((ROOT / "MEASURED_PEAKS.json").read_text())["hbm_gbs"]
    except Exception:
        return 6650.0


def report(rows, name, cfg, ms_list, bytes_moved):
    ms = statistics.median(ms_list)
    gbs = bytes_moved / ms / 1e6
    rows.append({"kernel": name, "config": cfg, "ms_median": round(ms, 4), "ms_min": round(min(ms_list), 4),
                 "GBps": round(gbs, 1), "frac_of_measured_hbm": round(gbs / peak_gbs(), 3)})
    print(json.dumps(rows[-1]), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--json", default=None)
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    rows = []
    it = dict(iters=5 if args.quick else 20, warmup=3)
    want = lambda k: (not args.only) or k in args.only.split(",")  # noqa: E731

    if want("lab1"):
        for n in ([1 << 25] if args.quick else [10 ** 6, 1 << 25, 1 << 28]):
            a = torch.randn(n, dtype=torch.float64, device=dev)
            b = torch.randn(n, dtype=torch.float64, device=dev)
            c = torch.empty_like(a)
            report(rows, "lab1.vecsub", f"n={n}", cuda_time_ms(lambda: ops.vecsub(a, b, c), **it), 24 * n)
            report(rows, "lab1.vecsub_geom<<<512,512>>>", f"n={n}",
                   cuda_time_ms(lambda: ops.vecsub_geom(a, b, 512, 512, c), **it), 24 * n)
            report(rows, "torch.sub (library, for scale)", f"n={n}",
                   cuda_time_ms(lambda: torch.sub(a, b, out=c), **it), 24 * n)
            del a, b, c

    if want("lab2"):
        for h, w in ([(16384, 16384)] if args.quick else [(1024, 1024), (16384, 16384), (32768, 32768)]):
            img = torch.randint(-2 ** 31, 2 ** 31 - 1, (h, w), dtype=torch.int32, device=dev)
            out = torch.empty_like(img)
            for v in ("stream", "tma", "cluster", "generic"):
                report(rows, f"lab2.roberts[{v}]", f"{w}x{h}",
                       cuda_time_ms(lambda: ops.roberts(img, out, variant=v), **it), 8 * h * w)
            report(rows, "lab2.roberts_geom block32x32 grid16x16", f"{w}x{h}",
                   cuda_time_ms(lambda: ops.roberts_geom(img, (32, 32), (16, 16), out), **it), 8 * h * w)
            report(rows, "torch copy_ (library, for scale)", f"{w}x{h}",
                   cuda_time_ms(lambda: out.copy_(img), **it), 8 * h * w)
            del img, out

    if want("lab3"):
        for h, w in ([(8192, 8192)] if args.quick else [(2048, 2048), (8192, 8192), (16384, 16384)]):
            img = torch.randint(-2 ** 31, 2 ** 31 - 1, (h, w), dtype=torch.int32, device=dev)
            rng = np.random.default_rng(0)
            classes = [np.stack([rng.integers(0, w, 4096), rng.integers(0, h, 4096)], axis=1) for _ in range(32)]
            params = ops.lab3_params_from_points(img, classes)
            scratch = torch.empty(1 << 24, dtype=torch.uint8, device=dev)
            for v in ("direct", "lut"):
                report(rows, f"lab3.classify[{v}] nc=32", f"{w}x{h}",
                       cuda_time_ms(lambda: ops.lab3_classify(img, params, variant=v, scratch=scratch),
                                    iters=3 if v == "direct" else it["iters"], warmup=2), 8 * h * w)
            report(rows, "lab3.classify_geom<<<256,256>>> nc=32", f"{w}x{h}",
                   cuda_time_ms(lambda: ops.lab3_classify_geom(img, params, 256, 256), iters=2, warmup=1), 8 * h * w)
            pts, off = ops.classify.pack_points(classes)
            d_pts, d_off = torch.from_numpy(pts).to(dev), torch.from_numpy(off).to(dev)
            mom = torch.zeros((32, 10), dtype=torch.int64, device=dev)
            report(rows, "lab3.moments+finalize nc=32 np=4096", f"{w}x{h}",
                   cuda_time_ms(lambda: ops.lab3_finalize(ops.lab3_moments(img, d_pts, d_off, 32, mom.zero_())), **it),
                   32 * 4096 * 12)
            del img

    if want("lab5"):
        for n in ([1 << 26] if args.quick else [1 << 20, 1 << 26, 1 << 28, 1 << 30]):
            for dt in (torch.int32, torch.float32, torch.uint8):
                if dt != torch.int32 and n > (1 << 28):
                    continue
                if dt == torch.uint8:
                    src = torch.randint(0, 256, (n,), dtype=torch.uint8, device=dev)
                elif dt == torch.int32:
                    src = torch.randint(-2 ** 31, 2 ** 31 - 1, (n,), dtype=torch.int32, device=dev)
                else:
                    src = torch.randint(-2 ** 31, 2 ** 31 - 1, (n,), dtype=torch.int32, device=dev).view(torch.float32)
                    src = torch.nan_to_num(src, nan=1.0)
                keys, tmp = torch.empty_like(src), torch.empty_like(src)
                scratch = torch.empty(ops.sort_scratch_bytes(n, dt), dtype=torch.uint8, device=dev)
                # time copy+sort, subtract the copy (the sort is in place)
                t_copy = statistics.median(cuda_time_ms(lambda: keys.copy_(src), **it))
                t_both = cuda_time_ms(lambda: (keys.copy_(src), ops.sort_(keys, tmp, scratch)), **it)
                es = src.element_size()
                passes = 9 if es == 4 else 2
                report(rows, f"lab5.sort[{str(dt).split('.')[-1]}]", f"n={n}", [t - t_copy for t in t_both], passes * es * n)
                if n <= (1 << 28) and dt != torch.uint8:
                    t_ref = cuda_time_ms(lambda: torch.sort(src), iters=3, warmup=1)
                    report(rows, f"torch.sort[{str(dt).split('.')[-1]}] (library, for scale)", f"n={n}", t_ref, passes * es * n)
                del src, keys, tmp, scratch

    print(f"libb200labs launches: {N.launch_count()}")
    if args.json:
        Path(args.json).write_text(json.dumps(rows, indent=1))


if __name__ == "__main__":
    main()
