"""The lab2 TMA kernel on an image whose byte offsets pass 2^32 (65536 x 20480 RGBA = 5 GiB in, 5 GiB
out): sampled 64-row blocks against the CPU oracle, including row 0, the first block beyond the
4 GiB line and the last rows (clamp-to-edge at the bottom), plus the same image filtered as two
bands with the halo pointer, which must give identical bytes."""
import pytest
import torch

from cuda_mpi_openmp_b200 import ops
from cuda_mpi_openmp_b200.utils import verify as V

pytestmark = pytest.mark.gpu


@pytest.mark.timeout(600)
def test_roberts_beyond_4gib_matches_oracle(cuda_device):
    free, _ = torch.cuda.mem_get_info(cuda_device)
    if free < 14 << 30:
        pytest.skip("needs 14 GiB of free device memory")
    w, h = 65536, 20480
    g = torch.Generator(device=cuda_device).manual_seed(11)
    img = torch.empty((h, w), dtype=torch.int32, device=cuda_device)
    for r0 in range(0, h, 4096):
        img[r0:r0 + 4096].copy_(torch.randint(-2 ** 31, 2 ** 31 - 1, (min(4096, h - r0), w), dtype=torch.int32,
                                              device=cuda_device, generator=g))
    img[9000:9004] = img[9000]                     # flat rows: zero gradient
    out = torch.empty_like(img)
    ops.roberts(img, out, variant="tma")
    starts = V.sample_block_starts(h, w, n_blocks=12, rows=64, seed=1) + [9000 - 32]
    assert any(4 * w * y >= 1 << 32 for y in starts) and starts[0] == 0 and h - 64 in starts
    res = V.roberts_blocks(img, out, None, starts, rows=64)
    assert res["ok"], res
    # two bands, the upper one reading its halo row through a pointer into the lower one
    split = 16384 + 37                             # the lower band starts beyond the 4 GiB line
    out2 = torch.empty_like(img)
    ops.roberts(img[:split], out2[:split], below_row=img[split], variant="tma")
    ops.roberts(img[split:], out2[split:], variant="tma")
    for r0 in range(0, h, 2048):
        assert torch.equal(out[r0:r0 + 2048], out2[r0:r0 + 2048]), r0
