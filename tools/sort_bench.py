#!/usr/bin/env python
"""lab5 radix sort: correctness + device time for one library configuration.

Library variants (another build via B200LABS_LIB=<path to .so>) are read once per process,
so each configuration is its own process:  python tools/sort_bench.py --n 26 --dist uniform

Sorting is timed out of place (b200_sort_ex reads `src`, writes `keys`), so no copy is
inside the timed region.  Every distribution is checked against torch.sort once."""
import argparse
import ctypes as C
import json
import os
import statistics
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from cuda_mpi_openmp_b200 import _native as N, ops  # noqa: E402
from cuda_mpi_openmp_b200.utils import cuda_time_ms  # noqa: E402

N.register("b200_sort_ex", C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int,
                                     C.c_void_p, C.c_size_t, C.c_void_p])


def make(dist: str, n: int, dtype, dev):
    g = torch.Generator(device=dev).manual_seed(1)
    if dist == "uniform":
        x = torch.randint(-2 ** 31, 2 ** 31 - 1, (n,), dtype=torch.int32, device=dev, generator=g)
    elif dist == "sorted":
        x = torch.arange(n, dtype=torch.int32, device=dev)
    elif dist == "reverse":
        x = torch.arange(n, 0, -1, dtype=torch.int32, device=dev)
    elif dist == "constant":
        x = torch.full((n,), 12345, dtype=torch.int32, device=dev)
    elif dist == "small":  # 1000 distinct values
        x = torch.randint(0, 1000, (n,), dtype=torch.int32, device=dev, generator=g)
    elif dist == "bits16":  # only 16 random bits, spread over the word
        x = torch.randint(0, 65536, (n,), dtype=torch.int32, device=dev, generator=g)
        x = (x & 0xFF) | ((x & 0xFF00) << 8)
    else:
        raise SystemExit(f"unknown distribution {dist}")
    if dtype == torch.float32:
        x = torch.nan_to_num(x.view(torch.float32), nan=1.0)
    return x


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", default="26", help="comma list of log2 sizes")
    ap.add_argument("--dist", default="uniform")
    ap.add_argument("--dtype", default="int32")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    dtype = {"int32": torch.int32, "float32": torch.float32}[args.dtype]
    tag = {"lib": os.environ.get("B200LABS_LIB", "default")}
    rows = []
    for lg in [int(v) for v in args.n.split(",")]:
        n = 1 << lg
        for dist in args.dist.split(","):
            src = make(dist, n, dtype, dev)
            keys, tmp = torch.empty_like(src), torch.empty_like(src)
            scratch = torch.empty(ops.sort_scratch_bytes(n, dtype), dtype=torch.uint8, device=dev)
            kt = ops.sort.KEY_TYPES[dtype]

            def run():
                N.check(N.lib().b200_sort_ex(N.ptr(src), N.ptr(keys), N.ptr(tmp), n, None, kt, N.ptr(scratch),
                                             scratch.numel(), N.current_stream_ptr()), "b200_sort_ex")

            run()
            torch.cuda.synchronize()
            ok = bool(torch.equal(keys, torch.sort(src).values)) if n <= (1 << 28) else bool(
                (keys[1:] >= keys[:-1]).all().item() and keys.sum(dtype=torch.int64) == src.sum(dtype=torch.int64))
            ms = cuda_time_ms(run, iters=args.iters, warmup=3)
            med = statistics.median(ms)
            row = {"n": f"2^{lg}", "dist": dist, "dtype": args.dtype, **tag, "ok": ok, "ms_median": round(med, 4),
                   "ms_min": round(min(ms), 4), "Gkeys_s": round(n / med / 1e6, 1),
                   "frac_of_hbm_roofline": round(36 * n / med / 1e6 / 6549.4, 3)}
            rows.append(row)
            print(json.dumps(row), flush=True)
            del src, keys, tmp, scratch
    if args.json:
        Path(args.json).write_text(json.dumps(rows, indent=1))
    if not all(r["ok"] for r in rows):
        raise SystemExit("MISMATCH")


if __name__ == "__main__":
    main()
