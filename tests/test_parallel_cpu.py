"""Host-side logic of the multi-GPU layer, on CPU: shard geometry, and the halo
hand-off order of the baseline path run by two gloo ranks against the CPU oracle."""
import os
import subprocess
import sys
import textwrap

import numpy as np

from cuda_mpi_openmp_b200.parallel import shard_rows
from cuda_mpi_openmp_b200.parallel.sharded import samples_per_rank_for


def test_shard_rows_cover_and_balance():
    for h in (1, 7, 8, 65536, 65537, 1003):
        for world in (1, 2, 4, 8):
            if h < world:
                continue
            bands = shard_rows(h, world)
            assert bands[0][0] == 0 and bands[-1][1] == h
            assert all(a1 == b0 for (_, a1), (b0, _) in zip(bands[:-1], bands[1:]))
            sizes = [b - a for a, b in bands]
            assert max(sizes) - min(sizes) <= 1 and min(sizes) >= 1


def test_samples_per_rank():
    for world in (1, 2, 4, 8):
        s = samples_per_rank_for(world)
        assert s * world == 8192


WORKER = textwrap.dedent('''
    import sys, numpy as np, torch, torch.distributed as dist
    sys.path.insert(0, sys.argv[1])
    from cuda_mpi_openmp_b200.models import RobertsFilter
    from cuda_mpi_openmp_b200.parallel import shard_rows
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    rng = np.random.default_rng(9)
    h, w = 41, 24
    img = rng.integers(0, 256, size=(h, w, 4), dtype=np.uint8)
    y0, y1 = shard_rows(h, world)[rank]
    band = torch.from_numpy(img[y0:y1].copy())
    halo = torch.empty((w, 4), dtype=torch.uint8)
    reqs = []
    if rank > 0:
        reqs.append(dist.isend(band[0].contiguous(), rank - 1))
    if rank < world - 1:
        reqs.append(dist.irecv(halo, rank + 1))
    for r in reqs:
        r.wait()
    # a band plus its halo row filtered as one image, the halo's output row dropped
    if rank < world - 1:
        ext = np.concatenate([band.numpy(), halo.numpy()[None]], axis=0)
        out = RobertsFilter.reference(ext)[:-1]
    else:
        out = RobertsFilter.reference(band.numpy())
    gathered = [None] * world
    dist.all_gather_object(gathered, out)
    if rank == 0:
        assert np.array_equal(np.concatenate(gathered, axis=0), RobertsFilter.reference(img))
        print("GLOO_OK")
    dist.destroy_process_group()
''')


def test_halo_handoff_two_gloo_ranks(built, tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", "29541", str(script), str(built)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300,
                       env={**os.environ, "CUDA_VISIBLE_DEVICES": ""})
    assert r.returncode == 0 and "GLOO_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


FD_WORKER = textwrap.dedent('''
    import os, sys, tempfile, torch, torch.distributed as dist
    sys.path.insert(0, sys.argv[1])
    from cuda_mpi_openmp_b200.parallel import DistPeerGroup
    from cuda_mpi_openmp_b200.parallel.vmm import exchange_fds
    dist.init_process_group("gloo")
    g = DistPeerGroup(device=torch.device("cpu"))
    # every rank hands each other rank two descriptors of files that name sender and receiver
    files, out = [], {}
    for dest in range(g.world):
        if dest == g.rank:
            continue
        fds = []
        for k in range(2):
            f = tempfile.TemporaryFile()
            f.write(b"from %d to %d #%d" % (g.rank, dest, k)); f.flush(); f.seek(0)
            files.append(f); fds.append(f.fileno())
        out[dest] = fds
    # a stray local process: right after the socket names went round, rank 0 also throws a
    # message with a wrong token (and a descriptor) plus a garbled one at every listener
    real_gather = dist.all_gather_object
    def gather_then_intrude(cards, card, group=None):
        real_gather(cards, card, group=group)
        if g.rank == 0:
            import socket
            for name, _pid, _tok in cards:
                for payload in (b"x" * 16 + b"   1   1", b"garbage"):
                    with socket.socket(socket.AF_UNIX, socket.SOCK_STREAM) as c:
                        c.connect(name)
                        socket.send_fds(c, [payload], [files[0].fileno()])
    dist.all_gather_object = gather_then_intrude
    got = exchange_fds(g, out)
    dist.all_gather_object = real_gather
    assert sorted(got) == [r for r in range(g.world) if r != g.rank], got
    assert all(len(v) == 2 for v in got.values()), got
    for sender, fds in got.items():
        for k, fd in enumerate(fds):
            with os.fdopen(fd, "rb") as f:
                f.seek(0)
                assert f.read() == b"from %d to %d #%d" % (sender, g.rank, k)
    dist.barrier()
    if g.rank == 0:
        print("FD_EXCHANGE_OK")
    dist.destroy_process_group()
''')


def test_fd_exchange_three_gloo_ranks(built, tmp_path):
    """The descriptor hand-off under the VMM/multicast allocator (parallel/vmm.py)."""
    script = tmp_path / "fd_worker.py"
    script.write_text(FD_WORKER)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=3",
           "--master-addr", "127.0.0.1", "--master-port", "29543", str(script), str(built)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300,
                       env={**os.environ, "CUDA_VISIBLE_DEVICES": ""})
    assert r.returncode == 0 and "FD_EXCHANGE_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
