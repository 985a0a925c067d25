"""The multi-rank kernels exercised on ONE GPU: `world` virtual ranks live in one
process, their "peer" pointers are each other's buffers, and the kernels run phase
by phase (push/signal for every rank first, wait/reduce afterwards) so that no
kernel ever waits on another launch."""
import numpy as np
import pytest
import torch

from cuda_mpi_openmp_b200 import _native as N
from cuda_mpi_openmp_b200 import ops
from cuda_mpi_openmp_b200.models import MahalanobisClassifier, RobertsFilter
from cuda_mpi_openmp_b200.parallel import LocalPeerGroup, shard_rows
from cuda_mpi_openmp_b200.parallel import sharded as S

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("variant", ["tma", "stream"])
def test_roberts_bands_in_symmetric_buffers(cuda_device, world, variant):
    rng = np.random.default_rng(world)
    h, w = 1000 + world, 768
    img = rng.integers(0, 256, size=(h, w, 4), dtype=np.uint8)
    group = LocalPeerGroup(world, cuda_device)
    bands = shard_rows(h, world)
    bufs = group.alloc(max(b - a for a, b in bands) * w * 4)
    for r, (a, b) in enumerate(bands):
        bufs[r].view(torch.uint8, b - a, w, 4).copy_(torch.from_numpy(img[a:b]))
    outs = []
    for r, (a, b) in enumerate(bands):
        below = bufs[r].peer_address(r + 1) if r + 1 < world else None   # "peer" pointer, raw address
        band = bufs[r].view(torch.int32, b - a, w)
        outs.append(ops.roberts(band, below_row=below, variant=variant).cpu().numpy())
    got = np.concatenate(outs, axis=0).view(np.uint8).reshape(h, w, 4)
    assert np.array_equal(got, RobertsFilter.reference(img))
    for b in bufs:
        b.free()


@pytest.mark.parametrize("world", [1, 2, 8])
def test_lab3_stats_allreduce_matches_single_gpu_bitwise(cuda_device, world):
    rng = np.random.default_rng(40 + world)
    h, w, nc = 257, 300, 32
    img = rng.integers(0, 256, size=(h, w, 4), dtype=np.uint8)
    classes = [np.stack([rng.integers(0, w, 3000), rng.integers(0, h, 3000)], axis=1) for _ in range(nc)]
    want = MahalanobisClassifier.reference_params(img, classes, mode="moments")
    pts, off = ops.classify.pack_points(classes)
    d_pts, d_off = torch.from_numpy(pts).to(cuda_device), torch.from_numpy(off).to(cuda_device)
    group = LocalPeerGroup(world, cuda_device)
    ws = group.workspaces()
    bands = shard_rows(h, world)
    d_bands = [torch.from_numpy(img[a:b].copy()).to(cuda_device) for a, b in bands]
    params = [torch.zeros((nc, 12), dtype=torch.float64, device=cuda_device) for _ in range(world)]
    for rep in range(3):                      # epochs alternate the slot parity
        epoch = group.next_epoch()
        for r, (a, b) in enumerate(bands):    # phase 1 on every rank: accumulate, push, signal
            S.fused_stats(d_bands[r], w, a, b - a, d_pts, d_off, nc, ws[r].ptrs_dev, r, world, epoch, params[r], None, 1)
        for r, (a, b) in enumerate(bands):    # phase 2: wait (already satisfied), reduce, finalize
            S.fused_stats(d_bands[r], w, a, b - a, d_pts, d_off, nc, ws[r].ptrs_dev, r, world, epoch, params[r], None, 2)
        for r in range(world):
            assert np.array_equal(params[r].cpu().numpy().view(np.int64), want.view(np.int64)), (rep, r)
    assert all(int(w_.local[group.ranks[0].ws_offset("error"):][:8].view(torch.int64).item()) == 0 for w_ in ws)


@pytest.mark.parametrize("world", [1, 2, 8])
def test_lab3_cooperative_lut_matches_oracle(cuda_device, world):
    rng = np.random.default_rng(60 + world)
    h, w, nc = 96, 128, 32
    img = rng.integers(0, 256, size=(h, w, 4), dtype=np.uint8)
    classes = [np.stack([rng.integers(0, w, 200), rng.integers(0, h, 200)], axis=1) for _ in range(nc)]
    params = MahalanobisClassifier.reference_params(img, classes, mode="moments")
    want = MahalanobisClassifier.reference_classify(img, params)
    d_params = torch.from_numpy(params).to(cuda_device)
    group = LocalPeerGroup(world, cuda_device)
    ws = group.workspaces()
    luts = group.alloc(1 << 24)
    epoch = group.next_epoch()
    for r in range(world):        # every rank classifies its slice of the colour cube into ALL tables
        S.lut_build_allgather(luts[r].peer_ptrs, nc, d_params, ws[r].ptrs_dev, r, world, epoch, 1)
    for r in range(world):
        S.lut_build_allgather(luts[r].peer_ptrs, nc, d_params, ws[r].ptrs_dev, r, world, epoch, 2)
    bands = shard_rows(h, world)
    for r, (a, b) in enumerate(bands):
        d = torch.from_numpy(img[a:b].copy()).to(cuda_device)
        S.apply_lut(d, luts[r].local_ptr)
        assert np.array_equal(d.cpu().numpy(), want[a:b]), r
    torch.cuda.synchronize()
    assert torch.equal(luts[0].local, luts[world - 1].local)
    # big band: the two-level image pass (colour-cell table in shared memory) against the one-level pass
    big = torch.randint(-2 ** 31, 2 ** 31 - 1, (2048, 2304), dtype=torch.int32, device=cuda_device)
    one, two = big.clone(), big.clone()
    coarse = torch.empty(int(N.lib().b200_lab3_coarse_bytes()), dtype=torch.uint8, device=cuda_device)
    S.apply_lut(one, luts[0].local_ptr)
    S.apply_lut(two, luts[world - 1].local_ptr, coarse)
    assert torch.equal(one, two)
    assert torch.equal((one.view(torch.uint8).reshape(2048, 2304, 4)[..., 3].long()),
                       luts[0].local.long()[(big & 0xFFFFFF).long()])
    for b in luts:
        b.free()


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_lab3_cooperative_table_with_cells_matches_oracle(cuda_device, world):
    """Second form of the cooperative build (colour table + cell table from one kernel, 16-byte
    stores, any world size): every rank's buffer ends up identical, the colour bytes equal the
    first form's table, the cell bytes equal what the separate coarsening pass derives, and the
    image pass on top of it matches the CPU classifier."""
    rng = np.random.default_rng(90 + world)
    h, w, nc = 96, 128, 32
    img = rng.integers(0, 256, size=(h, w, 4), dtype=np.uint8)
    classes = [np.stack([rng.integers(0, w, 200), rng.integers(0, h, 200)], axis=1) for _ in range(nc)]
    params = MahalanobisClassifier.reference_params(img, classes, mode="moments")
    want = MahalanobisClassifier.reference_classify(img, params)
    d_params = torch.from_numpy(params).to(cuda_device)
    group = LocalPeerGroup(world, cuda_device)
    ws = group.workspaces()
    nbytes = int(N.lib().b200_lab3_lut2_bytes())
    luts = group.alloc(nbytes)
    for rep in range(2):                       # second round: stale bytes of round one are overwritten
        epoch = group.next_epoch()
        for phase in (1, 2):
            for r in range(world):
                S.lut_build2_allgather(luts[r].peer_ptrs, 0, nc, d_params, ws[r].ptrs_dev, r, world, epoch, phase)
    torch.cuda.synchronize()
    for r in range(1, world):
        assert torch.equal(luts[0].local[:nbytes], luts[r].local[:nbytes]), r
    # reference table: the single-GPU older build + separate coarsening pass
    solo = LocalPeerGroup(1, cuda_device)
    old = solo.alloc(1 << 24)
    S.lut_build_allgather(old[0].peer_ptrs, nc, d_params, solo.workspaces()[0].ptrs_dev, 0, 1, solo.next_epoch(), 3)
    assert torch.equal(luts[0].local[: 1 << 24], old[0].local[: 1 << 24])
    big = torch.randint(-2 ** 31, 2 ** 31 - 1, (2048, 2304), dtype=torch.int32, device=cuda_device)
    coarse = torch.empty(int(N.lib().b200_lab3_coarse_bytes()), dtype=torch.uint8, device=cuda_device)
    one, two = big.clone(), big.clone()
    S.apply_lut(one, old[0].local_ptr, coarse)                 # derives `coarse` from the old table
    S.apply_lut2(two, luts[world - 1].local_ptr)
    torch.cuda.synchronize()
    assert torch.equal(coarse, luts[0].local[1 << 24: nbytes])
    assert torch.equal(one, two)
    for r, (a, b) in enumerate(shard_rows(h, world)):
        d = torch.from_numpy(img[a:b].copy()).to(cuda_device)
        S.apply_lut2(d, luts[r].local_ptr)
        assert np.array_equal(d.cpu().numpy(), want[a:b]), r
    for b in luts + old:
        b.free()


@pytest.mark.parametrize("world", [2, 3, 4, 6, 8])
@pytest.mark.parametrize("dtype,kind", [(np.int32, "random"), (np.float32, "random"), (np.int32, "dups"),
                                        (np.int32, "equal"), (np.uint32, "skewed")])
def test_sample_sort_fused_scatter(cuda_device, world, dtype, kind):
    rng = np.random.default_rng(50 + world)
    n_local = 100_000 + 37
    shards = []
    for r in range(world):
        if dtype == np.float32:
            k = (rng.standard_normal(n_local) * 1e3).astype(np.float32)
        else:
            info = np.iinfo(dtype)
            k = rng.integers(info.min, info.max, n_local, dtype=np.int64, endpoint=True).astype(dtype)
        if kind == "dups":
            k = k[rng.integers(0, 5, n_local)]
        elif kind == "equal":
            k[:] = 7
        elif kind == "skewed":                 # every rank holds a different key range
            k = (k // world + r * (info.max // world)).astype(dtype)
        shards.append(k)
    tdt = {np.int32: torch.int32, np.float32: torch.float32, np.uint32: torch.uint32}[dtype]
    group = LocalPeerGroup(world, cuda_device)
    ws = group.workspaces()
    capacity = int(n_local * 1.25) + 8192
    buckets = group.alloc(capacity * 4)
    d_keys = [torch.from_numpy(s.copy()).to(cuda_device) for s in shards]
    spr = S.samples_per_rank_for(world)
    off = {k: group.ranks[0].ws_offset(k) for k in ("sort_splitters", "sort_recv_count", "error")}
    e1, e2 = group.next_epoch(), group.next_epoch()
    for r in range(world):
        S.mgpu_splitters(d_keys[r], spr, ws[r].ptrs_dev, r, world, e1, 1)
    for r in range(world):
        S.mgpu_splitters(d_keys[r], spr, ws[r].ptrs_dev, r, world, e1, 2)
    split0 = ws[0].local[off["sort_splitters"]:][: 8 * world].view(torch.int64).cpu()
    for r in range(1, world):               # every rank derived the same splitters
        assert torch.equal(ws[r].local[off["sort_splitters"]:][: 8 * world].view(torch.int64).cpu(), split0)
    reserve = (world % 4 == 0)      # exercise both reservation schemes
    if reserve:                     # exact two-phase: count, one remote atomic per destination
        for r in range(world):
            S.mgpu_count(d_keys[r], ws[r].local_ptr + off["sort_splitters"], capacity, ws[r], r, world)
    for r in range(world):
        S.mgpu_scatter(d_keys[r], ws[r].local_ptr + off["sort_splitters"], buckets[r].peer_ptrs, None, capacity,
                       ws[r], r, world, e2, 1 | (8 if reserve else 0))
    for r in range(world):
        S.mgpu_scatter(d_keys[r], ws[r].local_ptr + off["sort_splitters"], buckets[r].peer_ptrs, None, capacity,
                       ws[r], r, world, e2, 2 | (8 if reserve else 0))
    parts, counts = [], []
    for r in range(world):
        bucket = buckets[r].view(tdt, capacity)
        tmp = torch.empty_like(bucket)
        scratch = torch.empty(ops.sort_scratch_bytes(capacity, tdt), dtype=torch.uint8, device=cuda_device)
        S.sort_devn(bucket, tmp, ws[r].local_ptr + off["sort_recv_count"], scratch)
        cnt = int(ws[r].local[off["sort_recv_count"]:][:8].view(torch.int64).item())
        assert int(ws[r].local[off["error"]:][:8].view(torch.int64).item()) == 0
        counts.append(cnt)
        parts.append(bucket[:cnt].cpu().numpy())
    assert sum(counts) == world * n_local
    assert max(counts) <= capacity and max(counts) < 1.15 * n_local + 8192     # balanced, duplicates included
    got = np.concatenate(parts)
    assert np.array_equal(got, np.sort(np.concatenate(shards)))
    for b in buckets:
        b.free()


def test_bucket_overflow_is_detected_not_fatal(cuda_device):
    """A destination bucket that is too small must raise the sticky error word and drop
    the excess keys — never write out of bounds, never hang."""
    world, n_local = 2, 50_000
    rng = np.random.default_rng(77)
    group = LocalPeerGroup(world, cuda_device)
    ws = group.workspaces()
    capacity = 20_000                                   # deliberately far too small
    buckets = group.alloc((capacity + 64) * 4)
    guard = 0x5A5A5A5A
    for b in buckets:
        b.view(torch.int32, capacity + 64)[capacity:] = guard
    d_keys = [torch.from_numpy(rng.integers(-2**31, 2**31 - 1, n_local, dtype=np.int64).astype(np.int32)).to(cuda_device)
              for _ in range(world)]
    off = {k: group.ranks[0].ws_offset(k) for k in ("sort_splitters", "sort_recv_count", "error")}
    e1, e2 = group.next_epoch(), group.next_epoch()
    for phase in (1, 2):
        for r in range(world):
            S.mgpu_splitters(d_keys[r], S.samples_per_rank_for(world), ws[r].ptrs_dev, r, world, e1, phase)
    for phase in (1, 2):
        for r in range(world):
            S.mgpu_scatter(d_keys[r], ws[r].local_ptr + off["sort_splitters"], buckets[r].peer_ptrs, None, capacity,
                           ws[r], r, world, e2, phase)
    torch.cuda.synchronize()
    errors = [int(w_.local[off["error"]:][:8].view(torch.int64).item()) for w_ in ws]
    assert any(e & 2 for e in errors)                   # kErrSortOverflow
    for r, b in enumerate(buckets):
        assert bool((b.view(torch.int32, capacity + 64)[capacity:] == guard).all()), r   # no overrun
        assert int(ws[r].local[off["sort_recv_count"]:][:8].view(torch.int64).item()) <= capacity
        b.free()
