"""Worker for the multi-process GPU test (launched by torchrun, one rank per GPU)."""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from cuda_mpi_openmp_b200.models import MahalanobisClassifier, RobertsFilter  # noqa: E402
from cuda_mpi_openmp_b200.parallel import ShardedClassifier, ShardedRoberts, ShardedSort, init_distributed  # noqa: E402


def main():
    import torch.distributed as dist

    g = init_distributed()
    rank, world, dev = g.rank, g.world, g.device
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
    for impl, variant in (("fused", "auto"), ("nccl", "auto"), ("fused", "lut"), ("nccl", "lut"), ("nvls", "lut")):
        try:
            m = ShardedClassifier(g, w, h, classes, impl=impl, variant=variant)
        except Exception as exc:  # noqa: BLE001 - NVLS needs NVSwitch multicast + torch symmetric memory
            if impl != "nvls":
                raise
            if rank == 0:
                print("NVLS path skipped:", type(exc).__name__, exc)
            continue
        assert m.coop_lut == (impl in ("fused", "nvls") and variant == "lut")
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
    g.barrier()
    if rank == 0:
        print("MP_WORKER_OK world", world)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
