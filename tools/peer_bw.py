#!/usr/bin/env python
"""What a kernel gets out of NVLink by access width (torchrun, >= 2 GPUs): every rank copies
256 MiB between its own memory and rank (r+1)'s buffer with 4-byte stores, 16-byte stores,
4-byte loads, 16-byte loads and 16 KB TMA bulk stores, all ranks at once (max over ranks).
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/peer_bw.py"""
import ctypes as C
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from cuda_mpi_openmp_b200 import _native as N  # noqa: E402
from cuda_mpi_openmp_b200.parallel import init_distributed  # noqa: E402

N.register("b200_selftest_peer_copy", C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p])


def main():
    g = init_distributed()
    rank, world, dev = g.rank, g.world, g.device
    nbytes = 256 << 20
    sym = g.alloc(nbytes)
    near = torch.empty(nbytes // 4, dtype=torch.int32, device=dev).fill_(rank + 1)
    far = sym.peer_ptrs[(rank + 1) % world]
    res = {"world": world, "bytes": nbytes}
    names = {0: "store_4B", 1: "store_16B", 2: "load_4B", 3: "load_16B", 4: "tma_bulk_store_16KB"}
    for mode, name in names.items():
        for blocks in (148 * 4, 148 * 16):
            def run():
                N.check(N.lib().b200_selftest_peer_copy(C.c_void_p(far), N.ptr(near), nbytes // 4, mode, blocks,
                                                        N.current_stream_ptr()), "peer_copy")
            run()
            g.barrier()
            ts = []
            for _ in range(5):
                g.barrier()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                run()
                b.record()
                torch.cuda.synchronize()
                ts.append(a.elapsed_time(b))
            ms = g.all_reduce_max(sorted(ts)[2])
            res[f"{name}_blocks{blocks}"] = {"ms": round(ms, 4), "GBps_per_gpu": round(nbytes / ms / 1e6, 1)}
    g.barrier()
    ok = bool((sym.view(torch.int32, nbytes // 4)[:1024] == ((rank - 1) % world) + 1).all().item())
    res["landed"] = bool(g.all_reduce_max(0.0 if ok else 1.0) == 0.0)
    if rank == 0:
        print(json.dumps(res, indent=1))
        Path("gpurun_out").mkdir(exist_ok=True)
        Path("gpurun_out/peer_bw.json").write_text(json.dumps(res, indent=1))
    g.barrier()
    sym.free()
    g.dist.destroy_process_group()


if __name__ == "__main__":
    main()
