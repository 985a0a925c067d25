"""Latency numbers of the peer-memory runtime, measured on the device (torchrun, >= 2 GPUs):

  * flag ping-pong rank 0 <-> rank 1 over a CUDA-IPC mapping and over a VMM mapping, with
    release/acquire and with relaxed system-scope accesses (globaltimer, 2000 round trips);
  * `b200_symm_barrier` (one launch: signal every peer, wait for every peer), CUDA events;
  * multimem.red over n uint64 (the NVLS add used by the lab3 statistics), CUDA events.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 \
        tools/symm_selftest.py --out gpurun_out/symm_selftest_n2.json
"""
import argparse
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from cuda_mpi_openmp_b200.parallel import init_distributed, vmm  # noqa: E402
from cuda_mpi_openmp_b200.utils.timing import cuda_time_ms  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    ap.add_argument("--iters", type=int, default=2000)
    args = ap.parse_args()
    g = init_distributed()
    rank, world, dev = g.rank, g.world, g.device
    res = {"world": world, "caps": vmm.query(), "granularity": vmm.granularity(world)}

    n = 1 << 16
    buf = vmm.alloc(g, n * 8)
    ipc = g.alloc(256)
    base = 0
    if world >= 2:
        for relaxed in (False, True):
            for name, mine, peers in (("vmm", buf.local_ptr, buf.peer_ptrs), ("ipc", ipc.local_ptr, ipc.peer_ptrs)):
                key = f"pingpong_{name}_{'relaxed' if relaxed else 'release_acquire'}_rtt_ns"
                if rank < 2:
                    ns = vmm.pingpong(mine, peers[1 - rank], rank == 0, args.iters, base, dev, relaxed=relaxed)
                    if rank == 0:
                        res[key] = round(ns, 1)
                base += args.iters
                g.barrier()
    buf.local.zero_()
    g.barrier()

    def timed(fn, reps=200):
        fn()
        g.barrier()
        ts = []
        for _ in range(5):
            g.barrier()
            ts.append(cuda_time_ms(lambda: [fn() for _ in range(reps)], warmup=0, iters=1)[0] / reps)
        return g.all_reduce_max(sorted(ts)[len(ts) // 2])

    res["symm_barrier_ms"] = round(timed(g.device_barrier), 5)
    if buf.mc_ptr:
        for cnt in (320, n):
            res[f"multimem_red_{cnt}_u64_ms"] = round(timed(lambda c=cnt: vmm.multimem(buf.mc_ptr, c, "red", value=1)), 5)
    g.barrier()
    if rank == 0:
        line = json.dumps(res)
        print(line)
        if args.out:
            Path(args.out).parent.mkdir(parents=True, exist_ok=True)
            Path(args.out).write_text(line + "\n")
    ipc.free()
    buf.free()
    g.dist.destroy_process_group()


if __name__ == "__main__":
    main()
