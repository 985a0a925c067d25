#!/usr/bin/env python
"""Tiny launcher for ncu: runs ONE lab kernel a few times at a representative size.
    python tools/profile_kernel.py lab2_tma|lab2_stream|lab1|lab3_direct|lab3_lut|lab5 [reps]"""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from cuda_mpi_openmp_b200 import ops  # noqa: E402


def main():
    what = sys.argv[1]
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    dev = torch.device("cuda:0")
    if what.startswith("lab2"):
        h = w = 16384
        if what.endswith("_big"):          # 4 GiB in + 4 GiB out, the flagship's row pitch
            h, w, what = 16384, 65536, what[:-4]
        img = torch.randint(-2 ** 31, 2 ** 31 - 1, (h, w), dtype=torch.int32, device=dev)
        out = torch.empty_like(img)
        for _ in range(reps):
            ops.roberts(img, out, variant=what.split("_")[1])
    elif what == "lab1":
        n = 1 << 25
        a = torch.randn(n, dtype=torch.float64, device=dev)
        b = torch.randn(n, dtype=torch.float64, device=dev)
        c = torch.empty_like(a)
        for _ in range(reps):
            ops.vecsub(a, b, c)
    elif what.startswith("lab3"):
        h = w = 8192
        img = torch.randint(-2 ** 31, 2 ** 31 - 1, (h, w), dtype=torch.int32, device=dev)
        rng = np.random.default_rng(0)
        classes = [np.stack([rng.integers(0, w, 4096), rng.integers(0, h, 4096)], axis=1) for _ in range(32)]
        params = ops.lab3_params_from_points(img, classes)
        scratch = torch.empty(1 << 24, dtype=torch.uint8, device=dev)
        for _ in range(reps):
            ops.lab3_classify(img, params, variant=what.split("_")[1], scratch=scratch)
    elif what == "lab5":
        n = 1 << 26
        src = torch.randint(-2 ** 31, 2 ** 31 - 1, (n,), dtype=torch.int32, device=dev)
        keys, tmp = torch.empty_like(src), torch.empty_like(src)
        scratch = torch.empty(ops.sort_scratch_bytes(n, torch.int32), dtype=torch.uint8, device=dev)
        for _ in range(reps):
            keys.copy_(src)
            ops.sort_(keys, tmp, scratch)
    else:
        raise SystemExit(f"unknown kernel {what}")
    torch.cuda.synchronize()
    print("done", what)


if __name__ == "__main__":
    main()
