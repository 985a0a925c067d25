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
(`ShardedRoberts.run_host`): every step copies the band from pinned host memory into
the symmetric device band, filters it with the same fused kernel (halo row read from
the neighbour GPU) and copies the result back.  `labs` carries the other labs' kernels
at the same N (fewer steps) for BASELINE.md's table.  `verify` holds output checks run
outside the timed regions (sampled blocks against the CPU oracles, sortedness and
checksums, bit-identical class parameters); any false verdict makes the exit code 3.
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
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extras", action="store_true")
    ap.add_argument("--e2e-steps", type=int, default=None)
    ap.add_argument("--ref-shape", default="stock",
                    help="reference arm: 'stock' = its fixed block 32x32 / grid 16x16; 'best' = the fastest "
                         "shape of tools/reference_sweep.py (profiles/reference_sweep.json); or 'bx,by,gx,gy'")
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


# ===================================================================== binaries, both arms
def write_synthetic_image(path, w, rows, seed):
    """`int w, int rows, RGBA[]` file of random pixels (a 1024-row random block, repeated)."""
    import numpy as np

    rng = np.random.default_rng(seed)
    with open(path, "wb") as f:
        f.write(np.array([w, rows], dtype=np.int32).tobytes())
        block = rng.integers(0, 256, size=min(rows, 1024) * w * 4, dtype=np.uint8)
        left = rows
        while left > 0:
            n = min(left, 1024)
            f.write(block[: n * w * 4].tobytes())
            left -= n


def binary_step(exe, stdin, env, n_chunks):
    """One step through a contract binary's stdin/file protocol: n_chunks invocations, each a fresh
    process (spawn, CUDA context, file read, H2D, kernel, D2H, file write).  Returns (sum of the
    binaries' own kernel ms, wall ms)."""
    kernel_ms, t0 = 0.0, time.perf_counter()
    for _ in range(n_chunks):
        r = subprocess.run([str(exe)], input=stdin, capture_output=True, text=True, env=env)
        if r.returncode != 0:
            raise RuntimeError(f"{exe} failed: {r.stderr[-500:]}")
        kernel_ms += float(r.stdout.split("<")[1].split(" ms>")[0])
    return kernel_ms, (time.perf_counter() - t0) * 1e3


def own_binary_e2e(args, group, rank, world, local_rank, w, h):
    """The like-for-like end-to-end arm: this repository's lab2/src/to_plot_cuda_exe driven exactly
    as bench.py drives the reference binary (same /dev/shm files, same stdin protocol, same
    <= 2^30-pixel chunks, a fresh process per chunk), plus one run per rank on its WHOLE band
    (64-bit sizes: the reference's int pixel index cannot do that)."""
    import numpy as np

    from cuda_mpi_openmp_b200.models import RobertsFilter

    exe = ROOT / "lab2" / "src" / "to_plot_cuda_exe"
    h_local = h // world
    rows_per_chunk = max(1, min(h_local, (1 << 30) // w))
    n_chunks = (h_local + rows_per_chunk - 1) // rows_per_chunk
    shm = Path("/dev/shm") if Path("/dev/shm").is_dir() else Path("/tmp")
    in_path, out_path = shm / f"b200own_in_{rank}.data", shm / f"b200own_out_{rank}.data"
    env = {**os.environ, "CUDA_VISIBLE_DEVICES": str(local_rank)}
    env.pop("B200LABS_WARM", None)
    res, ok = {}, True
    try:
        for label, rows, chunks, steps in (("chunked", rows_per_chunk, n_chunks, 2), ("whole_band", h_local, 1, 1)):
            if label == "whole_band" and n_chunks == 1:
                res[label] = "same as chunked (the band is one chunk)"
                continue
            # input + output file of every rank must fit the RAM disk: decided together, so that no
            # rank leaves the others waiting in a barrier
            import shutil
            need = 2 * (8 + 4 * w * rows) * world + (1 << 30)
            short = group.all_reduce_max(1.0 if shutil.disk_usage(shm).free < need else 0.0) > 0
            if short:
                res[label] = f"skipped: {shm} has less than {need >> 30} GiB free"
                continue
            write_synthetic_image(in_path, w, rows, seed=1234 + rank)
            stdin = f"-1\n-1\n-1\n-1\n{in_path}\n{out_path}\n"
            group.barrier()
            binary_step(exe, stdin, env, 1)                       # warm-up: page cache, driver files
            group.barrier()
            k_ms, wall = [], []
            for _ in range(steps):
                k, wl = binary_step(exe, stdin, env, chunks)
                k_ms.append(k)
                wall.append(wl)
            my_k, my_wall = group.all_reduce_max(sum(k_ms) / steps), group.all_reduce_max(sum(wall) / steps)
            total_bytes = BYTES_PER_PX * w * rows * chunks * world
            res[label] = {"value": round(total_bytes / my_wall / 1e6, 3), "unit": "GB/s", "ms_per_step": round(my_wall, 1),
                          "kernel_ms_per_step": round(my_k, 3), "steps": steps, "invocations_per_step": chunks,
                          "rows_per_invocation": rows}
            # the file the binary wrote, first and last 64 rows, against the CPU oracle
            with open(in_path, "rb") as f:
                head_in = np.frombuffer(f.read(8 + 65 * w * 4), dtype=np.uint8, offset=8).reshape(65, w, 4)
                f.seek(8 + (rows - 64) * w * 4)
                tail_in = np.frombuffer(f.read(64 * w * 4), dtype=np.uint8).reshape(64, w, 4)
            with open(out_path, "rb") as f:
                hdr = np.frombuffer(f.read(8), dtype=np.int32)
                head_out = np.frombuffer(f.read(64 * w * 4), dtype=np.uint8).reshape(64, w, 4)
                f.seek(8 + (rows - 64) * w * 4)
                tail_out = np.frombuffer(f.read(64 * w * 4), dtype=np.uint8).reshape(64, w, 4)
            size_ok = os.path.getsize(out_path) == 8 + rows * w * 4 and hdr.tolist() == [w, rows]
            ok = ok and size_ok and bool(np.array_equal(RobertsFilter.reference(head_in)[:64], head_out)) \
                and bool(np.array_equal(RobertsFilter.reference(tail_in), tail_out))
    finally:
        for p in (in_path, out_path):
            try:
                p.unlink()
            except OSError:
                pass
    ok = bool(group.all_reduce_max(0.0 if ok else 1.0) == 0.0)
    res["api"] = ("lab2/src/to_plot_cuda_exe through the reference's stdin/file protocol on /dev/shm: process start, "
                  "CUDA context, file read, H2D, kernel, D2H, file write, every invocation")
    return res, {"ok": ok, "checked": "first and last 64 rows of each output file vs CPU oracle, header, size"}


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
    with ClockSampler(gpu_index=local_rank, period_ms=1) as clocks:
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

    # ------------------------------------------------------------ verify the flagship output (untimed)
    verify = {}
    from cuda_mpi_openmp_b200.utils import verify as V
    below_t = None
    if world > 1 and rank < world - 1:                 # the neighbour's first row, read through the peer mapping
        below_t = model.sym_in.peer_view(rank + 1, torch.int32, w).clone()
    starts = V.sample_block_starts(model.h_local, w, n_blocks=16, rows=64, seed=rank)
    v2 = V.roberts_blocks(model.band, model.out, below_t, starts, rows=64)
    v2["byte_offsets_ge_2^32_sampled"] = bool(any(4 * w * y >= (1 << 32) for y in starts))
    v2["ok"] = bool(group.all_reduce_max(0.0 if v2["ok"] else 1.0) == 0.0)
    verify["lab2_blocks_vs_cpu_oracle"] = v2

    # ------------------------------------------------------------ end to end (host -> host)
    e2e = None
    if not args.no_e2e:
        e2e_steps = args.e2e_steps or max(2, min(args.steps, 5))
        host_in = torch.empty((model.h_local, w), dtype=torch.int32, pin_memory=True)
        host_out = torch.empty((model.h_local, w), dtype=torch.int32, pin_memory=True)
        host_in.copy_(model.band)                      # synthetic band, now resident in pinned host memory
        ref_out = model.out.clone()                    # device-path result of the same band
        torch.cuda.synchronize()
        group.barrier()
        e2e_ms = timed_steps(group, lambda: model.run_host(host_in, host_out), e2e_steps, 1)
        # the whole end-to-end output against the device-path output, compared on the device
        same, piece = True, max(1, (1 << 30) // (4 * w))
        for r0 in range(0, model.h_local, piece):
            up = host_out[r0:r0 + piece].to(dev, non_blocking=True)
            same = same and bool(torch.equal(up, ref_out[r0:r0 + piece]))
        del ref_out
        same = bool(group.all_reduce_max(0.0 if same else 1.0) == 0.0)
        verify["lab2_e2e_equals_device_path"] = {"ok": same, "rows_compared": model.h_local * world}
        e2e = {"value": round(total_bytes / e2e_ms / 1e6, 2), "unit": "GB/s", "ms_per_step": round(e2e_ms, 3),
               "steps": e2e_steps, "h2d_bytes_per_step": 4 * w * h, "d2h_bytes_per_step": 4 * w * h,
               "api": "ShardedRoberts.run_host (pinned host -> symmetric device band -> fused kernel with the "
                      "peer-memory halo -> pinned host; chunked, three streams)",
               "matches_device_path": same}
        del host_in, host_out
    flag_out_sample = model.out[:2, :8].cpu().tolist()
    group.barrier()
    model.free()
    del model
    torch.cuda.empty_cache()

    e2e_binary = None
    if not args.no_e2e and (ROOT / "lab2" / "src" / "to_plot_cuda_exe").exists():
        e2e_binary, verify["lab2_binary_output_vs_cpu_oracle"] = own_binary_e2e(args, group, rank, world, local_rank, w, h)

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
        verify["lab3"] = verify_lab3(group, m3, classes, s3)
        labs["lab3_stats_allreduce_32x4096pts"]["path"] = (
            "nvls (multimem.red)" if getattr(m3, "nvls", False) else "p2p stores" if args.exchange == "fused" else "nccl")
        if world > 1 and args.exchange == "fused":     # the other in-kernel all-reduce, for the record
            other = "p2p" if m3.nvls else "nvls"
            try:
                m3o = ShardedClassifier(group, s3, s3, classes, impl=other)
                m3o.band.copy_(m3.band)
                t = timed_steps(group, m3o.fit, 20, 3)
                same = bool(torch.equal(m3o.params.view(torch.int64), m3.params.view(torch.int64)))
                labs[f"lab3_stats_allreduce_{other}"] = {"ms": round(t, 4), "bit_identical": same}
                m3o.free()
            except Exception as exc:  # noqa: BLE001
                labs[f"lab3_stats_allreduce_{other}"] = {"unavailable": str(exc)[:200]}
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
        verify["lab5"] = verify_lab5(group, m5, n_total)
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
            "e2e": e2e, "e2e_binary": e2e_binary, "gpu_launches": int(launches_timed), "labs": labs,
            "verify": verify,
            "published_reference_note": "BASELINE.md has no published number for this config (reference numbers "
                                        "are A6000 launch-latency-bound <= 0.18 ms figures on <= 1 MP images)",
            "out_sample": flag_out_sample[0][:4],
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    if not all(v.get("ok", False) for v in verify.values()):
        if rank == 0:
            print(f"[bench] VERIFY FAILED: {json.dumps(verify)}", file=sys.stderr)
        sys.exit(3)


def verify_lab3(group, m3, classes, size):
    """lab3 outside the timed region: (a) the all-reduced class parameters are bit-identical to
    the CPU oracle's (exact integer moments of every band summed over the ranks, then the same
    finalize arithmetic); (b) >= 1 M pixels of the classified image, sampled as whole rows from
    every band, carry the class the CPU classifier assigns under those parameters."""
    import numpy as np
    import torch

    from cuda_mpi_openmp_b200 import _native as N
    from cuda_mpi_openmp_b200.models import MahalanobisClassifier
    from cuda_mpi_openmp_b200.ops import classify as CL

    world, rank = group.world, group.rank
    m3.step()
    m3.check()                                            # raises on an in-kernel peer time-out
    band = m3.band.cpu().numpy()                          # RGB untouched by classify, alpha = class
    pts, off = CL.pack_points(classes, w=size, h=size)
    mom = np.zeros((len(classes), CL.N_MOMENTS), dtype=np.int64)
    N.lib().b200_cpu_lab3_moments(N.ptr(band), size, m3.y0, m3.h_local, N.ptr(pts), N.ptr(off), len(classes),
                                  N.ptr(mom))
    mom_t = torch.from_numpy(mom).to(group.device)
    if world > 1:
        group.dist.all_reduce(mom_t, op=group.dist.ReduceOp.SUM, group=group.pg)
    mom = mom_t.cpu().numpy()
    params = np.zeros((len(classes), CL.N_PARAMS), dtype=np.float64)
    N.lib().b200_cpu_lab3_finalize(N.ptr(mom), len(classes), N.ptr(params))
    got = m3.params.cpu().numpy()
    params_ok = bool(np.array_equal(params.view(np.int64), got.view(np.int64)))
    rows = max(8, (64 + world - 1) // world)              # >= 64 rows x 16384 px = 1 M pixels in total
    ys = np.unique(np.concatenate([[0, m3.h_local - 1], np.random.default_rng(rank).integers(0, m3.h_local, rows)]))
    sample = np.ascontiguousarray(band[ys]).view(np.uint8).reshape(len(ys), size, 4)
    want = MahalanobisClassifier.reference_classify(sample, params)
    bad = int((want[..., 3] != sample[..., 3]).sum())
    ok = params_ok and bad == 0
    ok = bool(group.all_reduce_max(0.0 if ok else 1.0) == 0.0)
    return {"params_bit_identical_to_cpu_oracle": params_ok, "pixels_sampled": int(len(ys) * size * world),
            "pixels_wrong": bad, "ok": ok}


def verify_lab5(group, m5, n_total):
    """lab5 outside the timed region: every rank's bucket is sorted, bucket r's last key is not
    above bucket r+1's first key, the bucket sizes add up to n, and (sum, xor) of all keys is
    the same before and after (nothing lost, nothing invented)."""
    import torch

    from cuda_mpi_openmp_b200.utils import verify as V

    world = group.world
    m5.step()
    torch.cuda.synchronize()
    res = m5.result()
    cnt = int(res.numel())
    sorted_ok = V.is_sorted(res)
    s_in, x_in = V.checksum(m5.keys)
    s_out, x_out = V.checksum(res)
    first = int(res[0].item()) if cnt else 2 ** 31 - 1
    last = int(res[-1].item()) if cnt else -2 ** 31
    t = torch.tensor([cnt, s_in, s_out, x_in, x_out, first, last, int(sorted_ok)], dtype=torch.int64,
                     device=group.device)
    if world > 1:
        allt = [torch.empty_like(t) for _ in range(world)]
        group.dist.all_gather(allt, t, group=group.pg)
    else:
        allt = [t]
    rows = [x.tolist() for x in allt]
    total = sum(r[0] for r in rows)
    sums_ok = sum(r[1] for r in rows) == sum(r[2] for r in rows)
    xi = xo = 0
    for r in rows:
        xi ^= r[3]
        xo ^= r[4]
    nonempty = [r for r in rows if r[0] > 0]
    bounds_ok = all(a[6] <= b[5] for a, b in zip(nonempty, nonempty[1:]))
    all_sorted = all(r[7] == 1 for r in rows)
    err = group.ws_error() if world > 1 else 0
    ok = all_sorted and bounds_ok and total == n_total and sums_ok and xi == xo and err == 0
    return {"buckets_sorted": all_sorted, "bucket_boundaries_ordered": bounds_ok, "keys_total": total,
            "sum_checksum_equal": sums_ok, "xor_checksum_equal": xi == xo, "error_word": err,
            "bucket_sizes": [r[0] for r in rows], "ok": ok}


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
    write_synthetic_image(in_path, w, rows_per_chunk, seed=1234 + rank)
    env = {**os.environ, "CUDA_VISIBLE_DEVICES": str(local_rank)}
    shape = (32, 32, 16, 16)
    if args.ref_shape == "best":
        try:
            sweep = json.loads((ROOT / "profiles" / "reference_sweep.json").read_text())
            best = next(v for k, v in sweep.items() if k.startswith("65536"))["reference_best_shape"]
            shape = (*best["block"], *best["grid"])
        except Exception:  # noqa: BLE001 - no sweep on record: stay with the stock shape
            args.ref_shape = "stock"
    elif args.ref_shape != "stock":
        shape = tuple(int(v) for v in args.ref_shape.split(","))
    stdin = "%d\n%d\n%d\n%d\n" % shape + f"{in_path}\n{out_path}\n"

    def one_step():
        return binary_step(exe, stdin, env, n_chunks)

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
                           "launch": "block %dx%d, grid %dx%d (%s)" % (*shape, "the reference's own fixed shape"
                                                                       if args.ref_shape == "stock" else
                                                                       "--ref-shape " + args.ref_shape),
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
