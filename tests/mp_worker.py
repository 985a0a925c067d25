"""Worker for the multi-process GPU test (launched by torchrun, one rank per GPU)."""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from cuda_mpi_openmp_b200.models import MahalanobisClassifier, RobertsFilter  # noqa: E402
from cuda_mpi_openmp_b200.parallel import ShardedClassifier, ShardedRoberts, ShardedSort, init_distributed  # noqa: E402


def runtime_selftests(g):
    """VMM / multicast runtime across real processes: descriptors over unix sockets, peer
    mappings, flag ping-pong over both allocators, multimem red / ld_reduce / st."""
    from cuda_mpi_openmp_b200.parallel import vmm

    rank, world, dev = g.rank, g.world, g.device
    n = 2048
    buf = vmm.alloc(g, n * 8)
    loc = buf.view(torch.int64, n)
    # unicast: every rank writes its row of rank (rank+1)'s buffer through the peer mapping
    nxt = (rank + 1) % world
    peer = vmm._wrap(buf.peer_ptrs[nxt], buf.nbytes, dev).view(torch.int64)
    peer[rank * 16:(rank + 1) * 16] = 1000 + rank
    g.barrier()
    prv = (rank - 1) % world
    assert torch.all(loc[prv * 16:(prv + 1) * 16] == 1000 + prv), "vmm peer store did not land"
    g.barrier()
    loc.zero_()
    g.barrier()
    # ping-pong between rank 0 and rank 1: one flag on each side, stored through the peer mapping
    lat = {}
    ipc = g.alloc(256)                                          # the cudaMalloc + CUDA-IPC allocator
    if rank < 2 and world >= 2:
        other = 1 - rank
        for name, mine, theirs in (("vmm", buf.local_ptr + 8 * (n - 1), buf.peer_ptrs[other] + 8 * (n - 1)),
                                   ("ipc", ipc.local_ptr, ipc.peer_ptrs[other])):
            lat[name] = vmm.pingpong(mine, theirs, rank == 0, 2000, 0, dev)
            lat[name + "_relaxed"] = vmm.pingpong(mine, theirs, rank == 0, 2000, 2000, dev, relaxed=True)
    g.barrier()
    ipc.free()
    if rank == 0 and lat:
        print("PINGPONG_OK round trip ns:", {k: round(v) for k, v in lat.items()})
    loc.zero_()
    g.barrier()
    if buf.mc_ptr:
        idx = np.arange(1, n + 1, dtype=np.int64)
        vmm.multimem(buf.mc_ptr, n, "red", value=rank + 1)
        g.barrier()
        total = world * (world + 1) // 2
        assert np.array_equal(loc.cpu().numpy(), total * idx), "multimem.red sum wrong"
        g.barrier()
        loc.copy_(torch.from_numpy((rank + 1) * 10 * idx).to(dev))
        g.barrier()
        out = torch.zeros(n, dtype=torch.int64, device=dev)
        vmm.multimem(buf.mc_ptr, n, "ld_reduce", out=out)
        torch.cuda.synchronize()
        assert np.array_equal(out.cpu().numpy(), total * 10 * idx), "multimem.ld_reduce sum wrong"
        g.barrier()
        if rank == world - 1:
            vmm.multimem(buf.mc_ptr, n, "st", value=500)
        g.barrier()
        assert np.array_equal(loc.cpu().numpy(), 499 + idx), "multimem.st broadcast wrong"
        if rank == 0:
            print("MULTIMEM_OK (red, ld_reduce, st over %d GPUs)" % world)
    elif rank == 0:
        print("multicast not available: multimem self-tests skipped")
    del loc, peer
    buf.free()


def main():
    import torch.distributed as dist

    g = init_distributed()
    rank, world, dev = g.rank, g.world, g.device
    if world > 1:
        runtime_selftests(g)
    rng = np.random.default_rng(123)                       # same stream on every rank
    h, w = 64 * world + 3, 512
    img = rng.integers(0, 256, size=(h, w, 4), dtype=np.uint8)
    want = RobertsFilter.reference(img)
    for impl in ("fused", "nccl"):
        m = ShardedRoberts(g, w, h, impl=impl, variant="tma")
        m.band.copy_(torch.from_numpy(img[m.y0:m.y1].copy()).view(torch.int32).reshape(m.h_local, w))
        g.barrier()
        for _ in range(2):
            out = m.step()
        got = out.cpu().numpy().view(np.uint8).reshape(m.h_local, w, 4)
        assert np.array_equal(got, want[m.y0:m.y1]), f"roberts {impl} rank {rank}"
        g.barrier()

    nc = 32
    classes = [np.stack([rng.integers(0, w, 500), rng.integers(0, h, 500)], axis=1) for _ in range(nc)]
    p_want = MahalanobisClassifier.reference_params(img, classes, mode="moments")
    c_want = MahalanobisClassifier.reference_classify(img, p_want)
    for impl, variant in (("fused", "auto"), ("nccl", "auto"), ("fused", "lut"), ("p2p", "lut"), ("nccl", "lut"),
                          ("nvls", "lut")):
        try:
            m = ShardedClassifier(g, w, h, classes, impl=impl, variant=variant)
        except RuntimeError as exc:                      # NVLS needs NVSwitch multicast
            if impl != "nvls" or "not available" not in str(exc):
                raise
            if rank == 0:
                print("NVLS path skipped:", type(exc).__name__, exc)
            continue
        assert m.coop_lut == (impl in ("fused", "nvls", "p2p") and variant == "lut")
        assert not (impl == "p2p" and m.nvls)
        for rep in range(3):
            m.band.copy_(torch.from_numpy(img[m.y0:m.y1].copy()).view(torch.int32).reshape(m.h_local, w))
            g.barrier()
            out = m.step()
            assert np.array_equal(m.params.cpu().numpy().view(np.int64), p_want.view(np.int64)), f"lab3 params {impl}"
            got = out.cpu().numpy().view(np.uint8).reshape(m.h_local, w, 4)
            assert np.array_equal(got, c_want[m.y0:m.y1]), f"lab3 classify {impl}/{variant} rank {rank}"
        assert g.ws_error() == 0
        if impl == "nvls" and rank == 0:
            print("NVLS_STATS_OK (multimem.red all-reduce, bit-identical parameters)")
        g.barrier()
        m.free()

    n_local = 300_000 + 11
    for impl in ("fused", "nccl"):
        for dt, npdt in ((torch.int32, np.int32), (torch.float32, np.float32)):
            m = ShardedSort(g, n_local, dt, impl=impl)
            for rep in range(2):
                krng = np.random.default_rng(1000 * rep + rank)
                k = (krng.standard_normal(n_local) * 1e5).astype(npdt)
                m.keys.copy_(torch.from_numpy(k))
                g.barrier()
                m.step()
                torch.cuda.synchronize()
                assert g.ws_error() == 0, f"sort {impl} error word {g.ws_error()}"
                mine = m.result().cpu()
                gathered = [None] * world
                dist.all_gather_object(gathered, (k, mine.numpy()))
                if rank == 0:
                    allk = np.sort(np.concatenate([a for a, _ in gathered]))
                    got = np.concatenate([b for _, b in gathered])
                    assert np.array_equal(got, allk), f"sort {impl} {dt}"
                    sizes = [b.size for _, b in gathered]
                    assert max(sizes) <= m.capacity
    # ---- adversarial inputs for the fused sort: nothing may be lost, whatever the distribution
    def check_sorted_multiset(m, k, tag):
        g.barrier()
        m.step()
        mine = m.result().cpu().numpy()                  # collective: settles (and retries) the step
        assert g.ws_error() == 0, f"{tag}: error word {g.ws_error()}"
        gathered = [None] * world
        dist.all_gather_object(gathered, (k, mine))
        if rank == 0:
            allk = np.sort(np.concatenate([a for a, _ in gathered]))
            got = np.concatenate([b for _, b in gathered])
            assert got.size == allk.size, f"{tag}: {allk.size - got.size} keys lost"
            assert np.array_equal(got, allk), tag

    n_adv = 200_000 + 7
    m = ShardedSort(g, n_adv, torch.int32)
    cases = {
        "all_equal": np.full(n_adv, 42, dtype=np.int32),
        "one_ranks_range": (np.random.default_rng(rank).integers(0, 1000, n_adv)).astype(np.int32),
        "presorted_shards": np.sort(np.random.default_rng(7 + rank).integers(-2**31, 2**31 - 1, n_adv)).astype(np.int32),
        "disjoint_ranges": (np.random.default_rng(rank).integers(0, 1 << 20, n_adv) + (rank << 24)).astype(np.int32),
    }
    for tag, k in cases.items():
        m.keys.copy_(torch.from_numpy(k))
        check_sorted_multiset(m, k, tag)
    assert m.retries == 0, "the unique-composite splitters keep every one of these balanced"
    g.barrier()
    m.free()
    # buckets far too small: the step overflows, result() grows them on every rank and re-runs
    old_slack, ShardedSort.SLACK = ShardedSort.SLACK, 0.4
    try:
        m = ShardedSort(g, n_adv, torch.int32)
        k = np.random.default_rng(99 + rank).integers(-2**31, 2**31 - 1, n_adv).astype(np.int32)
        m.keys.copy_(torch.from_numpy(k))
        check_sorted_multiset(m, k, "forced_overflow")
        assert m.retries >= 1 and m.capacity >= n_adv * 0.8
        check_sorted_multiset(m, k, "after_growth")      # the grown buckets keep working
        g.barrier()
        m.free()
    finally:
        ShardedSort.SLACK = old_slack
    if rank == 0:
        print("SORT_ADVERSARIAL_OK (all equal, one range, presorted, disjoint, forced overflow -> retry)")
    g.barrier()
    if rank == 0:
        print("MP_WORKER_OK world", world)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
