"""The output checks bench.py runs outside its timed regions (utils/verify.py), on CPU tensors."""
import numpy as np
import torch

from cuda_mpi_openmp_b200.models import RobertsFilter
from cuda_mpi_openmp_b200.utils import verify as V


def test_xor_and_sum_checksums():
    rng = np.random.default_rng(3)
    for n in (1, 2, 3, 7, 1000, 4097):
        a = rng.integers(-2 ** 31, 2 ** 31 - 1, n, dtype=np.int64).astype(np.int32)
        want = int(np.bitwise_xor.reduce(a.view(np.uint32)))
        t = torch.from_numpy(a)
        assert V.xor_reduce(t) == want
        s, x = V.checksum(t)
        assert s == int(a.astype(np.int64).sum()) and x == want
        perm = torch.from_numpy(rng.permutation(a))
        assert V.checksum(perm) == (s, x)
    assert V.xor_reduce(torch.zeros(0, dtype=torch.int32)) == 0


def test_is_sorted_in_pieces():
    t = torch.arange(10_000, dtype=torch.int32)
    assert V.is_sorted(t, chunk=777) and V.is_sorted(t[:1]) and V.is_sorted(t[:0])
    u = t.clone()
    u[777] = 5                      # the break sits exactly on a piece boundary
    assert not V.is_sorted(u, chunk=777)


def test_block_starts_cover_first_last_and_the_4gib_line():
    ys = V.sample_block_starts(h=65536, w=65536, n_blocks=16, rows=64)
    assert ys[0] == 0 and ys[-1] == 65536 - 64 and len(ys) == 16
    assert any(4 * 65536 * y >= 1 << 32 for y in ys)
    small = V.sample_block_starts(h=100, w=512, n_blocks=16, rows=64)
    assert small[0] == 0 and small[-1] == 36


def test_roberts_blocks_catch_a_wrong_row_and_use_the_halo(lib):
    rng = np.random.default_rng(5)
    img = rng.integers(0, 256, size=(200, 96, 4), dtype=np.uint8)
    want = RobertsFilter.reference(img)
    as_i32 = lambda a: torch.from_numpy(a.copy()).view(torch.int32).reshape(a.shape[0], a.shape[1])  # noqa: E731
    # the band is the first 120 rows; the row below it is the image's row 120
    band, out, below = as_i32(img[:120]), as_i32(want[:120]), as_i32(img[120:121])[0]
    ys = V.sample_block_starts(120, 96, n_blocks=4, rows=32)
    res = V.roberts_blocks(band, out, below, ys, rows=32)
    assert res["ok"] and res["halo_block_checked"]
    assert not V.roberts_blocks(band, out, None, ys, rows=32)["ok"]      # clamping there is wrong
    out[5, 7] ^= 0x0100
    assert V.roberts_blocks(band, out, below, ys, rows=32)["bad_blocks"] == [0]
