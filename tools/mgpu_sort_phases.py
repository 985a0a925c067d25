#!/usr/bin/env python
"""Per-phase device times of the fused multi-GPU sort (run under torchrun):
splitters / count+reserve / partition+scatter / local sort.  Max over ranks."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from cuda_mpi_openmp_b200.parallel import ShardedSort, init_distributed  # noqa: E402
from cuda_mpi_openmp_b200.parallel import sharded as S  # noqa: E402


def main():
    g = init_distributed()
    world, r = g.world, g.rank
    n_total = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 31
    m = ShardedSort(g, n_total // world, torch.int32, impl="fused")
    gen = torch.Generator(device=g.device).manual_seed(7 + r)
    for o in range(0, m.n_local, 1 << 28):
        part = m.keys[o:o + (1 << 28)]
        part.copy_(torch.randint(-2 ** 31, 2 ** 31 - 1, (part.numel(),), dtype=torch.int32, device=g.device, generator=gen))
    ws_local = m.ws.local_ptr
    split = ws_local + m._off["sort_splitters"]

    def timed(fn):
        g.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        return g.all_reduce_max(a.elapsed_time(b))

    for rep in range(3):
        t1 = timed(lambda: S.mgpu_splitters(m.keys, m.samples, m.ws.ptrs_dev, r, world, g.next_epoch(), 3))
        t2 = timed(lambda: S.mgpu_count(m.keys, split, m.capacity, m.ws, r, world)) if m.reserve else 0.0
        t3 = timed(lambda: S.mgpu_scatter(m.keys, split, m.sym_bucket.peer_ptrs, None, m.capacity, m.ws, r, world,
                                          g.next_epoch(), 1 | 2 | (8 if m.reserve else 0)))
        t4 = timed(lambda: S.sort_devn(m.bucket, m.tmp, ws_local + m._off["sort_recv_count"], m.scratch))
        m._settled = False
        cnt, err = m.count(), g.ws_error()          # count() is collective: every rank calls it
        if r == 0:
            print(f"rep {rep}: splitters {t1:.3f} ms, count+reserve {t2:.3f} ms, scatter {t3:.3f} ms, "
                  f"local sort {t4:.3f} ms, bucket {cnt} keys, err {err}", flush=True)
    g.barrier()
    torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
