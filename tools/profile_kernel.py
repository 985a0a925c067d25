#!/usr/bin/env python
"""Tiny launcher for ncu: runs ONE lab kernel a few times at a representative size.
    python tools/profile_kernel.py lab2_tma|lab2_stream|lab1|lab3_direct|lab3_lut|lab5|mgpu [reps]
`mgpu` runs the multi-rank kernels as 8 virtual ranks on ONE GPU, phase by phase (so the
"remote" stores land in local HBM: it shows the compute side of the fused kernels)."""
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
        scratch = torch.empty((1 << 24) + (1 << 17), dtype=torch.uint8, device=dev)
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
    elif what == "mgpu":
        from cuda_mpi_openmp_b200.parallel import LocalPeerGroup, ShardedClassifier, SoloPeerGroup
        from cuda_mpi_openmp_b200.parallel import sharded as S

        def ms(fn):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            return a.elapsed_time(b)

        # lab3 as a world of one: fused statistics kernel, hoisted table build, table apply
        h = w = 8192
        rng = np.random.default_rng(0)
        classes = [np.stack([rng.integers(0, w, 4096), rng.integers(0, h, 4096)], axis=1) for _ in range(32)]
        m3 = ShardedClassifier(SoloPeerGroup(dev), w, h, classes, impl="fused", variant="lut")
        m3.band.copy_(torch.randint(-2 ** 31, 2 ** 31 - 1, (h, w), dtype=torch.int32, device=dev))
        for _ in range(reps):
            t = ms(m3.step)
        print(f"lab3 step {h}x{w} nc=32 world 1: {t:.3f} ms")
        # lab5: 8 virtual ranks x 2^24 keys, rank 0's partition+scatter is the kernel of interest
        world, n_local = 8, 1 << 24
        group = LocalPeerGroup(world, dev)
        ws = group.workspaces()
        capacity = int(n_local * 1.25) + 8192
        buckets = group.alloc(capacity * 4)
        keys = [torch.randint(-2 ** 31, 2 ** 31 - 1, (n_local,), dtype=torch.int32, device=dev) for _ in range(world)]
        spr = S.samples_per_rank_for(world)
        off = group.ranks[0].ws_offset("sort_splitters")
        for rep in range(reps):
            e1, e2 = group.next_epoch(), group.next_epoch()
            for ph in (1, 2):
                for r in range(world):
                    S.mgpu_splitters(keys[r], spr, ws[r].ptrs_dev, r, world, e1, ph)
            ts = [ms(lambda r=r: S.mgpu_scatter(keys[r], ws[r].local_ptr + off, buckets[r].peer_ptrs, None, capacity,
                                                ws[r], r, world, e2, 1)) for r in range(world)]
            for r in range(world):
                S.mgpu_scatter(keys[r], ws[r].local_ptr + off, buckets[r].peer_ptrs, None, capacity, ws[r], r, world,
                               e2, 2)
        print(f"lab5 partition+scatter, 2^24 keys/rank, 8 virtual ranks: {sorted(ts)[len(ts) // 2]:.3f} ms per rank "
              f"({n_local / sorted(ts)[len(ts) // 2] / 1e6:.1f} Gkeys/s)")
    else:
        raise SystemExit(f"unknown kernel {what}")
    torch.cuda.synchronize()
    print("done", what)


if __name__ == "__main__":
    main()
