#!/usr/bin/env python
"""Headline benchmark (driver contract).

Flagship = BASELINE.json's scaling config: the lab2 Roberts-cross filter on a synthetic
65536 x 65536 RGBA image, row-sharded over N B200s with the halo row pulled from the
neighbour GPU INSIDE the stencil kernel.  Metric: effective GB/s = 8 B/px (4 read + 4
written) x pixels / device time, whole job, strong scaling (total work fixed).

    python bench.py --gpus N --steps K --warmup W            # this framework
    python bench.py --impl reference --gpus N ...            # the unmodified reference

N > 1 is launched by the driver through torch.distributed.run (one rank per GPU); the
ranks time the same K steps with CUDA events, the maximum over ranks is reported by
rank 0 as ONE JSON line.  `e2e` is the same metric through the public pipeline API
(`RobertsFilter.run_host`): every step copies the band from pinned host memory to the
GPU, filters it and copies the result back.  `labs` carries the other labs' kernels
at the same N (fewer steps) for BASELINE.md's table.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "lab2_roberts_effective_GBps"
BYTES_PER_PX = 8


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["own", "reference"], default="own")
    ap.add_argument("--size", type=int, default=65536, help="image edge in pixels (w = h)")
    ap.add_argument("--variant", default="auto", help="lab2 kernel variant (auto|tma|stream|generic)")
    ap.add_argument("--exchange", choices=["fused", "nccl"], default="fused")
    ap.add_argument("--lab3-nvls", action="store_true", help="extras: also time the NVLS (multimem) statistics")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extras", action="store_true")
    ap.add_argument("--e2e-steps", type=int, default=None)
    ap.add_argument("--ref-max-seconds", type=float, default=420.0,
                    help="reference arm: cap on total wall time; steps are reduced (and reported) to fit")
    return ap.parse_args()


def env_world():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def peak_gbs():
    try:
        return json.loads((ROOT / "MEASURED_PEAKS.json").read_text())["hbm_gbs"], "measured"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback"


# =============================================================================== own arm
def fill_random_(t, seed):
    """Random bits into an int32 (or int32-viewable) device tensor, in 1 GiB pieces."""
    import torch

    g = torch.Generator(device=t.device).manual_seed(seed)
    flat = t.view(-1).view(torch.int32)
    chunk = 1 << 28
    for o in range(0, flat.numel(), chunk):
        part = flat[o:o + chunk]
        part.copy_(torch.randint(-2 ** 31, 2 ** 31 - 1, (part.numel(),), dtype=torch.int32, device=t.device,
                                 generator=g))


def timed_steps(group, step_fn, steps, warmup):
    """K steps between barrier+synchronize brackets, device time by CUDA events,
    maximum over ranks.  Returns ms per step."""
    import torch

    for _ in range(warmup):
        step_fn()
    group.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        step_fn()
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b)
    group.barrier()
    return group.all_reduce_max(ms) / steps


def run_own(args):
    import torch

    from cuda_mpi_openmp_b200 import _native as N
    from cuda_mpi_openmp_b200.models import RobertsFilter
    from cuda_mpi_openmp_b200.parallel import (ShardedClassifier, ShardedRoberts, ShardedSort, SoloPeerGroup,
                                                init_distributed)
    from cuda_mpi_openmp_b200.utils import ClockSampler

    rank, world, local_rank = env_world()
    if world > 1:
        # stdout carries exactly one JSON line: NCCL's own banner ("NCCL version ...", printed
        # whenever NCCL_DEBUG is set in the environment) goes to stderr instead
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        group = init_distributed()
    else:
        torch.cuda.set_device(0)
        group = SoloPeerGroup("cuda:0")
    if world != args.gpus and rank == 0:
        print(f"[bench] note: --gpus {args.gpus} but WORLD_SIZE={world}; using {world}", file=sys.stderr)
    dev = group.device
    w = h = args.size
    N.lib()

    # ------------------------------------------------------------ flagship (device-timed)
    model = ShardedRoberts(group, w, h, impl=args.exchange, variant=args.variant)
    fill_random_(model.band, seed=1234 + rank)
    group.barrier()
    launches0 = N.launch_count()
    with ClockSampler(gpu_index=local_rank) as clocks:
        ms = timed_steps(group, model.step, args.steps, args.warmup)
    launches_timed = (N.launch_count() - launches0) * args.steps // (args.steps + args.warmup)
    clock_summary = clocks.summary()
    # context: a plain device-to-device copy of the same band (same bytes read + written)
    copy_ms = timed_steps(group, lambda: model.out.copy_(model.band), 5, 2)
    model.step()                                       # model.out holds the filter output again
    torch.cuda.synchronize()
    total_bytes = BYTES_PER_PX * w * h
    value = total_bytes / ms / 1e6
    peak, peak_kind = peak_gbs()

    # ------------------------------------------------------------ end to end (host -> host)
    e2e = None
    if not args.no_e2e:
        e2e_steps = args.e2e_steps or max(2, min(args.steps, 5))
        host_in = torch.empty((model.h_local, w), dtype=torch.int32, pin_memory=True)
        host_out = torch.empty((model.h_local, w), dtype=torch.int32, pin_memory=True)
        host_in.copy_(model.band)                      # synthetic band, now resident in pinned host memory
        below_host = None
        if world > 1:                                  # halo row of the next band, exchanged once at set-up
            import torch.distributed as dist
            first = model.band[0].clone()
            halo = torch.empty_like(first)
            ops = []
            if rank > 0:
                ops.append(dist.P2POp(dist.isend, first, rank - 1))
            if rank < world - 1:
                ops.append(dist.P2POp(dist.irecv, halo, rank + 1))
            for req in dist.batch_isend_irecv(ops):
                req.wait()
            if rank < world - 1:
                below_host = halo.cpu().pin_memory()
        pipe = RobertsFilter(dev, variant=args.variant)
        e2e_ms = timed_steps(group, lambda: pipe.run_host(host_in, host_out, below_host), e2e_steps, 1)
        check = bool(torch.equal(host_out[:64], model.out[:64].cpu()))
        e2e = {"value": round(total_bytes / e2e_ms / 1e6, 2), "unit": "GB/s", "ms_per_step": round(e2e_ms, 3),
               "steps": e2e_steps, "h2d_bytes_per_step": 4 * w * h + (world - 1) * 4 * w,
               "d2h_bytes_per_step": 4 * w * h, "api": "RobertsFilter.run_host (pinned host -> GPU -> pinned host)",
               "matches_device_path": check}
        del host_in, host_out, pipe
    flag_out_sample = model.out[:2, :8].cpu().tolist()
    group.barrier()
    model.free()
    del model
    torch.cuda.empty_cache()

    # ------------------------------------------------------------ the other labs at this N
    labs = {}
    if not args.no_extras:
        import numpy as np
        from cuda_mpi_openmp_b200 import ops

        xs, xw = 5, 2
        # lab1: n = 2^28 doubles in total (6.4 GB moved), sliced over the ranks: no communication
        from cuda_mpi_openmp_b200.parallel import ShardedVectorSub
        n1 = 1 << 28
        m1 = ShardedVectorSub(group, n1)
        t = timed_steps(group, m1.step, 20, 3)
        labs["lab1_vecsub_n2^28"] = {"ms": round(t, 4), "GBps": round(24 * n1 / t / 1e6, 1),
                                     "frac_hbm": round(24 * n1 / t / 1e6 / peak / world, 3)}
        del m1
        if world == 1:
            n = 1 << 25
            a = torch.randn(n, dtype=torch.float64, device=dev)
            b = torch.randn(n, dtype=torch.float64, device=dev)
            c = torch.empty_like(a)
            t = timed_steps(group, lambda: ops.vecsub(a, b, c), 20, 3)
            labs["lab1_vecsub_n2^25"] = {"ms": round(t, 4), "GBps": round(24 * n / t / 1e6, 1),
                                         "frac_hbm": round(24 * n / t / 1e6 / peak, 3)}
            del a, b, c
        # lab3: 16384^2, 32 classes, statistics all-reduced in-kernel, then classify
        s3 = 16384
        rng = np.random.default_rng(7)
        classes = [np.stack([rng.integers(0, s3, 4096), rng.integers(0, s3, 4096)], axis=1) for _ in range(32)]
        m3 = ShardedClassifier(group, s3, s3, classes, impl=args.exchange)
        fill_random_(m3.band, seed=99 + rank)
        t = timed_steps(group, m3.step, xs, xw)
        labs["lab3_classify_16384^2_nc32"] = {"ms": round(t, 4), "GBps": round(8 * s3 * s3 / t / 1e6, 1),
                                              "frac_hbm": round(8 * s3 * s3 / t / 1e6 / peak / world, 3),
                                              "stats": "in-kernel all-reduce" if args.exchange == "fused" else "nccl"}
        t = timed_steps(group, m3.fit, 20, 3)
        labs["lab3_stats_allreduce_32x4096pts"] = {"ms": round(t, 4)}
        if args.lab3_nvls and world > 1:
            try:
                m3n = ShardedClassifier(group, s3, s3, classes, impl="nvls")
                m3n.band.copy_(m3.band)
                t = timed_steps(group, m3n.fit, 20, 3)
                same = bool(torch.equal(m3n.params.view(torch.int64), m3.params.view(torch.int64)))
                labs["lab3_stats_allreduce_nvls"] = {"ms": round(t, 4), "bit_identical_to_p2p": same}
                m3n.free()
            except Exception as exc:  # noqa: BLE001
                labs["lab3_stats_allreduce_nvls"] = {"unavailable": str(exc)[:200]}
        group.barrier()
        m3.free()
        del m3
        torch.cuda.empty_cache()
        # lab5: 2^31 32-bit keys over the ranks (sample sort, fused all-to-all)
        n_total = 1 << 31
        m5 = ShardedSort(group, n_total // world, torch.int32, impl=args.exchange)
        fill_random_(m5.keys, seed=555 + rank)
        t = timed_steps(group, m5.step, 3, 1)
        err = group.ws_error() if world > 1 else 0
        labs["lab5_sort_2^31_int32"] = {"ms": round(t, 3), "Gkeys_per_s": round(n_total / t / 1e6, 2),
                                        "bucket_keys_rank0": m5.count(), "error_word": err}
        group.barrier()
        m5.free()
        del m5
        torch.cuda.empty_cache()

    if rank == 0:
        line = {
            "metric": METRIC, "value": round(value, 2), "unit": "GB/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms, 4), "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "uint8 RGBA in/out, fp32 arithmetic (bit-exact to the reference kernel)",
            "data": "synthetic (random RGBA generated on the device; no dataset or checkpoint involved)",
            "impl": "own",
            "config": {"model": "lab2 Roberts-cross filter", "image": f"{w}x{h}", "global_batch": 1,
                       "seq_len": w * h, "parallelism": f"row-sharded x{world}, halo row by {args.exchange}",
                       "kernel_variant": args.variant,
                       "l2_policy": "inputs larger than L2 (>= 2 GiB read + 2 GiB written per GPU per step)"},
            "frac_of_hbm_roofline": round(value / (peak * world), 3), "roofline_peak_gbs_per_gpu": peak,
            "roofline_kind": f"of {peak_kind}",
            "context": {"torch_copy_same_band_GBps": round(BYTES_PER_PX * w * h / copy_ms / 1e6, 1),
                        "note": "library copy of the same band, max over ranks; footprints this large copy "
                                "slower than the 2+2 GiB copy MEASURED_PEAKS.json was taken on"},
            "clocks": {"sm_mhz": clock_summary["sm_mhz"], "sm_max_mhz": clock_summary["sm_max_mhz"],
                       "reasons": clock_summary["reasons"], "samples": clock_summary["samples"]},
            "e2e": e2e, "gpu_launches": int(launches_timed), "labs": labs,
            "published_reference_note": "BASELINE.md has no published number for this config (reference numbers "
                                        "are A6000 launch-latency-bound <= 0.18 ms figures on <= 1 MP images)",
            "out_sample": flag_out_sample[0][:4],
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


# ========================================================================= reference arm
def run_reference(args):
    """The unmodified reference binary (baseline/_ref/bin/lab2_to_plot, built from the
    untouched lab2/src/to_plot.cu) driven through its own stdin/file protocol at its own
    stock launch shape (block 32x32, grid 16x16).  Its `int` pixel indexing limits one
    invocation to < 2^31 px, so each rank's band is processed as chunks of <= 2^30 px."""
    rank, world, local_rank = env_world()
    exe = ROOT / "baseline" / "_ref" / "bin" / "lab2_to_plot"
    if not exe.exists():
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable":
                              "baseline/_ref/bin/lab2_to_plot missing (run baseline/install_reference.sh)"}))
        return
    import numpy as np

    dist = None
    if world > 1:
        import torch.distributed as dist  # gloo: keeps this process off the GPU
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("gloo")
    w = h = args.size
    h_local = h // world
    rows_per_chunk = max(1, min(h_local, (1 << 30) // w))
    n_chunks = (h_local + rows_per_chunk - 1) // rows_per_chunk
    shm = Path("/dev/shm") if Path("/dev/shm").is_dir() else Path("/tmp")
    in_path, out_path = shm / f"b200ref_in_{rank}.data", shm / f"b200ref_out_{rank}.data"
    rng = np.random.default_rng(1234 + rank)
    with open(in_path, "wb") as f:
        f.write(np.array([w, rows_per_chunk], dtype=np.int32).tobytes())
        block = rng.integers(0, 256, size=min(rows_per_chunk, 1024) * w * 4, dtype=np.uint8)
        left = rows_per_chunk
        while left > 0:
            n = min(left, 1024)
            f.write(block[: n * w * 4].tobytes())
            left -= n
    env = {**os.environ, "CUDA_VISIBLE_DEVICES": str(local_rank)}
    stdin = f"32\n32\n16\n16\n{in_path}\n{out_path}\n"

    def one_step():
        kernel_ms, t0 = 0.0, time.perf_counter()
        for _ in range(n_chunks):
            r = subprocess.run([str(exe)], input=stdin, capture_output=True, text=True, env=env)
            if r.returncode != 0:
                raise RuntimeError(f"reference binary failed: {r.stderr[-500:]}")
            kernel_ms += float(r.stdout.split("<")[1].split(" ms>")[0])
        return kernel_ms, (time.perf_counter() - t0) * 1e3

    def barrier():
        if dist is not None:
            dist.barrier()

    try:
        barrier()
        try:
            k0, wall0 = one_step()                    # first warm-up step doubles as a cost probe
        except Exception as exc:  # noqa: BLE001 - e.g. no GPU on this box
            if rank == 0:
                print(json.dumps({"impl": "reference", "unavailable": " ".join(str(exc).split())[:300]}))
            return
        budget_steps = max(1, int(args.ref_max_seconds * 1e3 / max(wall0, 1.0)) - 1)
        warmup = max(1, min(args.warmup, max(1, budget_steps // 4)))
        steps = max(1, min(args.steps, budget_steps - warmup + 1))
        if dist is not None:                          # every rank must run the same counts
            import torch
            t = torch.tensor([warmup, steps])
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
            warmup, steps = int(t[0]), int(t[1])
        for _ in range(warmup - 1):
            one_step()
        barrier()
        k_ms, wall_ms = [], []
        for _ in range(steps):
            k, wl = one_step()
            k_ms.append(k)
            wall_ms.append(wl)
        barrier()
        my_k, my_wall = sum(k_ms) / steps, sum(wall_ms) / steps
        if dist is not None:
            import torch
            t = torch.tensor([my_k, my_wall], dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            my_k, my_wall = float(t[0]), float(t[1])
        total_bytes = BYTES_PER_PX * w * h_local * world
        if rank == 0:
            print(json.dumps({
                "metric": METRIC, "value": round(total_bytes / my_k / 1e6, 2), "unit": "GB/s", "n_gpus": world,
                "steps": steps, "warmup": warmup, "ms_per_step": round(my_k, 4), "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "uint8 RGBA, fp32 arithmetic",
                "data": "synthetic (random RGBA written to /dev/shm)", "impl": "reference",
                "config": {"model": "lab2 Roberts-cross filter (reference lab2/src/to_plot.cu, unmodified, "
                                    "nvcc -arch=sm_100a)", "image": f"{w}x{h}", "global_batch": 1, "seq_len": w * h,
                           "parallelism": f"{world} independent processes, one band each, no halo exchange "
                                          "(the reference has no multi-GPU code)",
                           "launch": "block 32x32, grid 16x16 (the reference's own fixed shape)",
                           "chunks_per_rank": n_chunks, "rows_per_chunk": rows_per_chunk,
                           "note": "device ms = the binary's own cudaEvent kernel time summed over chunks; "
                                   "chunking needed because the reference indexes pixels with int"},
                "e2e": {"value": round(total_bytes / my_wall / 1e6, 3), "unit": "GB/s",
                        "ms_per_step": round(my_wall, 1),
                        "h2d_bytes_per_step": 4 * w * h_local * world, "d2h_bytes_per_step": 4 * w * h_local * world,
                        "api": "reference binary end to end (file read, H2D, kernel, D2H, file write)"},
                "gpu_launches": steps * n_chunks * world,
                "requested": {"steps": args.steps, "warmup": args.warmup},
            }), flush=True)
    finally:
        for p in (in_path, out_path):
            try:
                p.unlink()
            except OSError:
                pass
        if dist is not None:
            dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_own(args)


if __name__ == "__main__":
    main()
