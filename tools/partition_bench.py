#!/usr/bin/env python
"""Device time of the fused partition + scatter kernel (csrc/mgpu_lab5.cu) on ONE GPU: eight
virtual ranks whose buckets all live on this GPU, so the number is the kernel's own partition
cost with every "remote" store and atomic staying local (the NVLink part comes on top).
Usage: python tools/partition_bench.py [--lg 26] [--world 8]"""
import argparse
import json
import statistics
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from cuda_mpi_openmp_b200.parallel import LocalPeerGroup  # noqa: E402
from cuda_mpi_openmp_b200.parallel import sharded as S  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--lg", type=int, default=26)
    ap.add_argument("--world", type=int, default=8)
    ap.add_argument("--iters", type=int, default=8)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    n, world = 1 << a.lg, a.world
    g = LocalPeerGroup(world, dev)
    ws = g.workspaces()
    gen = torch.Generator(device=dev).manual_seed(3)
    keys = torch.randint(-2 ** 31, 2 ** 31 - 1, (n,), dtype=torch.int32, device=dev, generator=gen)
    small = [keys[: 1 << 16].clone() for _ in range(world)]          # the other ranks only feed the splitters
    capacity = n                                                       # rank 0's keys spread over 8 buckets
    buckets = g.alloc(capacity * 4)
    off = g.ranks[0].ws_offset("sort_splitters")
    spr = S.samples_per_rank_for(world)

    def prepare():
        e = g.next_epoch()
        for phase in (1, 2):
            for r in range(world):
                S.mgpu_splitters(keys if r == 0 else small[r], spr, ws[r].ptrs_dev, r, world, e, phase)

    times = []
    for it in range(a.iters + 2):
        prepare()
        e = g.next_epoch()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        t0.record()
        S.mgpu_scatter(keys, ws[0].local_ptr + off, buckets[0].peer_ptrs, None, capacity, ws[0], 0, world, e, 1)
        t1.record()
        torch.cuda.synchronize()
        if it >= 2:
            times.append(t0.elapsed_time(t1))
    # every key arrived somewhere: cursors add up to n, and the error word is clean
    cur = g.ranks[0].ws_offset("sort_cursor")
    total = sum(int(w.local[cur: cur + 8].view(torch.int64).item()) for w in ws)
    err = int(ws[0].local[g.ranks[0].ws_offset("error"):][:8].view(torch.int64).item())
    ms = statistics.median(times)
    print(json.dumps({"n_local": f"2^{a.lg}", "world": world, "ms_median": round(ms, 4), "Gkeys_s": round(n / ms / 1e6, 1),
                      "keys_delivered": total, "ok": total == n and err == 0}))


if __name__ == "__main__":
    main()
