"""The fused partition+scatter kernel with SEVERAL tiles per CTA.  The kernel is persistent
(2 CTAs per SM walk the tiles); the other multi-rank tests use ~10^5 keys per rank, i.e. at
most one tile per CTA, so the reuse of the shared-memory staging buffers from one tile to the
next is only exercised here: 4 virtual ranks x 6 M keys = 733 tiles per rank over 296 CTAs.
Checked exactly: every bucket sorted, globally ordered, and the multiset of all keys intact.
(Collected late on purpose: it is the slowest single-GPU multi-rank test.)"""
import numpy as np
import pytest
import torch

from cuda_mpi_openmp_b200 import ops
from cuda_mpi_openmp_b200.parallel import LocalPeerGroup
from cuda_mpi_openmp_b200.parallel import sharded as S

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [np.int32, np.float32])
def test_partition_many_tiles_per_cta(cuda_device, dtype):
    world, n_local = 4, 6_000_000 + 13
    rng = np.random.default_rng(2024)
    if dtype == np.float32:
        shards = [(rng.standard_normal(n_local) * 1e3).astype(np.float32) for _ in range(world)]
        tdt = torch.float32
    else:
        shards = [rng.integers(-2 ** 31, 2 ** 31 - 1, n_local, dtype=np.int64).astype(np.int32) for _ in range(world)]
        tdt = torch.int32
    group = LocalPeerGroup(world, cuda_device)
    ws = group.workspaces()
    capacity = int(n_local * 1.25) + 8192
    buckets = group.alloc(capacity * 4)
    d_keys = [torch.from_numpy(s).to(cuda_device) for s in shards]
    spr = S.samples_per_rank_for(world)
    off = {k: group.ranks[0].ws_offset(k) for k in ("sort_splitters", "sort_recv_count", "error")}
    e1, e2 = group.next_epoch(), group.next_epoch()
    for phase in (1, 2):
        for r in range(world):
            S.mgpu_splitters(d_keys[r], spr, ws[r].ptrs_dev, r, world, e1, phase)
    for phase in (1, 2):
        for r in range(world):
            S.mgpu_scatter(d_keys[r], ws[r].local_ptr + off["sort_splitters"], buckets[r].peer_ptrs, None, capacity,
                           ws[r], r, world, e2, phase)
    torch.cuda.synchronize()
    parts = []
    for r in range(world):
        assert int(ws[r].local[off["error"]:][:8].view(torch.int64).item()) == 0
        cnt = int(ws[r].local[off["sort_recv_count"]:][:8].view(torch.int64).item())
        assert 0 < cnt <= capacity
        bucket = buckets[r].view(tdt, capacity)
        # the partition alone: what landed in the bucket, before the local sort
        parts.append(bucket[:cnt].cpu().numpy().copy())
        tmp = torch.empty_like(bucket)
        scratch = torch.empty(ops.sort_scratch_bytes(capacity, tdt), dtype=torch.uint8, device=cuda_device)
        S.sort_devn(bucket, tmp, ws[r].local_ptr + off["sort_recv_count"], scratch)
        got_sorted = bucket[:cnt].cpu().numpy()
        assert np.array_equal(got_sorted, np.sort(parts[-1])), f"local sort of bucket {r}"
        parts[-1] = got_sorted
    assert sum(p.size for p in parts) == world * n_local
    for r in range(world - 1):                       # key ranges of consecutive buckets do not overlap
        assert parts[r][-1] <= parts[r + 1][0]
    assert max(p.size for p in parts) < 1.15 * n_local + 8192
    assert np.array_equal(np.concatenate(parts), np.sort(np.concatenate(shards)))
    for b in buckets:
        b.free()
