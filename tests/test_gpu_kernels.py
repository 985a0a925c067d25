"""GPU numerics: every kernel against (a) the bit-exact C++ oracle and (b) a plain
PyTorch restatement of the same op."""
import numpy as np
import pytest
import torch

from cuda_mpi_openmp_b200 import _native as N
from cuda_mpi_openmp_b200 import ops
from cuda_mpi_openmp_b200.models import MahalanobisClassifier, RadixSorter, RobertsFilter, VectorSub
from cuda_mpi_openmp_b200.utils import ImgData

pytestmark = pytest.mark.gpu


# ------------------------------------------------------------------------ lab1
@pytest.mark.parametrize("n", [1, 3, 4, 1000, 4099, (1 << 20) + 3, 1 << 24])
def test_vecsub_matches_torch(cuda_device, n):
    g = torch.Generator(device="cpu").manual_seed(n)
    a = (torch.rand(n, generator=g, dtype=torch.float64) - 0.5) * 2e100
    b = (torch.rand(n, generator=g, dtype=torch.float64) - 0.5) * 2e100
    before = N.launch_count()
    c = ops.vecsub(a.to(cuda_device), b.to(cuda_device))
    assert N.launch_count() > before
    assert torch.equal(c.cpu(), a - b)


def test_vecsub_unaligned_and_geom(cuda_device):
    a = torch.randn(5001, dtype=torch.float64, device=cuda_device)
    b = torch.randn(5001, dtype=torch.float64, device=cuda_device)
    assert torch.equal(ops.vecsub(a[1:], b[1:]), a[1:] - b[1:])      # 8-byte aligned only
    for grid, block in [(1, 32), (512, 512), (1024, 1024), (7, 33)]:
        assert torch.equal(ops.vecsub_geom(a, b, grid, block), a - b)
    with pytest.raises(N.NativeError):
        ops.vecsub_geom(a, b, 1, 2048)                                # invalid launch is reported


# ------------------------------------------------------------------------ lab2
def roberts_torch(img_u8: torch.Tensor) -> torch.Tensor:
    """fp32 PyTorch restatement (no control over FMA contraction: +-1 LSB)."""
    f = img_u8.to(torch.float32)
    y = 0.299 * f[..., 0] + 0.587 * f[..., 1] + 0.114 * f[..., 2]
    yp = torch.nn.functional.pad(y[None, None], (0, 1, 0, 1), mode="replicate")[0, 0]
    gx = yp[1:, 1:] - yp[:-1, :-1]
    gy = yp[:-1, 1:] - yp[1:, :-1]
    g = torch.sqrt(gx * gx + gy * gy).clamp(0, 255).to(torch.uint8)
    return torch.stack([g, g, g, img_u8[..., 3]], dim=-1)


SHAPES = {
    "generic": [(1, 1), (3, 3), (5, 7), (33, 130), (257, 255)],
    "stream": [(1, 4), (2, 128), (64, 132), (300, 512), (1000, 1028), (9, 4096)],
    # widths with 0 < w % 256 <= 128 make the right-half warps skip the last tile column:
    # with enough tile rows a CTA sees stored / skipped / stored (the staging-buffer case)
    "tma": [(1, 256), (33, 256), (300, 512), (1000, 1028), (2049, 2052), (31, 4096), (9500, 384),
            (4800, 1920)],
    "cluster": [(1, 256), (33, 256), (300, 512), (1000, 1028), (2049, 2052), (31, 4096), (70, 768),
                (9500, 384), (4800, 1920)],
}


@pytest.mark.timeout(180)
@pytest.mark.parametrize("variant", ["generic", "stream", "tma", "cluster", "auto"])
def test_roberts_bit_exact(cuda_device, variant):
    rng = np.random.default_rng(7)
    shapes = SHAPES.get(variant) or (SHAPES["generic"] + SHAPES["stream"])
    for h, w in shapes:
        img = rng.integers(0, 256, size=(h, w, 4), dtype=np.uint8)
        if h > 8:
            img[h // 3: h // 3 + 4] = img[h // 3]          # flat rows: exercises s == 0
        d = torch.from_numpy(img).to(cuda_device)
        out = ops.roberts(d, variant=variant)
        want = RobertsFilter.reference(img)
        assert np.array_equal(out.cpu().numpy(), want), (variant, h, w)
        t = roberts_torch(d)
        assert (out.to(torch.int16) - t.to(torch.int16)).abs().max().item() <= 1


@pytest.mark.timeout(180)
@pytest.mark.parametrize("variant", ["generic", "stream", "tma", "cluster"])
def test_roberts_row_sharded_with_halo_pointer(cuda_device, variant):
    """Row bands filtered independently, each reading its halo row through a pointer
    into the next band (what the multi-GPU path does with a peer pointer)."""
    rng = np.random.default_rng(11)
    h, w = 777, 1024
    img = rng.integers(0, 256, size=(h, w, 4), dtype=np.uint8)
    want = RobertsFilter.reference(img)
    cuts = [0, 200, 201, 555, h]
    bands = [torch.from_numpy(img[a:b].copy()).to(cuda_device) for a, b in zip(cuts[:-1], cuts[1:])]
    outs = []
    for i, band in enumerate(bands):
        below = bands[i + 1][0] if i + 1 < len(bands) else None
        outs.append(ops.roberts(band, below_row=below, variant=variant).cpu().numpy())
    assert np.array_equal(np.concatenate(outs, axis=0), want)


def test_roberts_goldens_and_real_images(cuda_device, repo_root, tmp_path):
    for name in ("test_01", "test_02"):
        src = ImgData(str(repo_root / "lab2/data" / f"{name}.txt"), out_dir=str(tmp_path), materialize=False).rgba
        gold = ImgData(str(repo_root / "lab2/data_out_gt" / f"{name}.txt"), out_dir=str(tmp_path), materialize=False).rgba
        out = ops.roberts(torch.from_numpy(src.copy()).to(cuda_device))
        assert np.array_equal(out.cpu().numpy(), gold)
    for name in ("lenna.data", "world_map.data"):
        src = ImgData(str(repo_root / "lab2/test_data" / name), materialize=False).rgba
        d = torch.from_numpy(src.copy()).to(cuda_device)
        want = RobertsFilter.reference(src)
        for variant in ("auto", "generic") + (("tma", "stream") if src.shape[1] % 4 == 0 else ()):
            assert np.array_equal(ops.roberts(d, variant=variant).cpu().numpy(), want), (name, variant)


def test_roberts_geom_sweep(cuda_device):
    rng = np.random.default_rng(3)
    img = rng.integers(0, 256, size=(97, 203, 4), dtype=np.uint8)
    d = torch.from_numpy(img).to(cuda_device)
    want = RobertsFilter.reference(img)
    for block, grid in [((32, 32), (16, 16)), ((16, 16), (32, 32)), ((2, 2), (16, 16)), ((1, 1), (1, 1))]:
        assert np.array_equal(ops.roberts_geom(d, block, grid).cpu().numpy(), want)
    with pytest.raises(N.NativeError):
        ops.roberts_geom(d, (512, 512), (1024, 1024))     # the README's invalid 512x512 block


def test_roberts_host_pipeline(cuda_device):
    rng = np.random.default_rng(5)
    img = rng.integers(0, 256, size=(1500, 512, 4), dtype=np.uint8)
    host = torch.from_numpy(img).view(torch.int32).reshape(1500, 512).pin_memory()
    model = RobertsFilter(cuda_device, chunk_bytes=512 * 4 * 200)     # 8 chunks over 3 streams
    out = model.run_host(host)
    assert np.array_equal(out.numpy().view(np.uint8).reshape(1500, 512, 4), RobertsFilter.reference(img))


# ------------------------------------------------------------------------ lab3
def random_classes(rng, w, h, nc, np_):
    return [np.stack([rng.integers(0, w, np_), rng.integers(0, h, np_)], axis=1) for _ in range(nc)]


def test_lab3_moments_and_params_bit_exact(cuda_device):
    rng = np.random.default_rng(21)
    h, w = 300, 257
    img = rng.integers(0, 256, size=(h, w, 4), dtype=np.uint8)
    classes = random_classes(rng, w, h, 32, 5000)
    d = torch.from_numpy(img).to(cuda_device)
    pts, off = ops.classify.pack_points(classes)
    d_pts, d_off = torch.from_numpy(pts).to(cuda_device), torch.from_numpy(off).to(cuda_device)
    mom = ops.lab3_moments(d, d_pts, d_off, 32)
    want_mom = np.zeros((32, 10), dtype=np.int64)
    N.lib().b200_cpu_lab3_moments(N.ptr(img), w, 0, h, N.ptr(pts), N.ptr(off), 32, N.ptr(want_mom))
    assert np.array_equal(mom.cpu().numpy(), want_mom)
    # banded accumulation (what row-sharded GPUs do) gives the same integers
    mom2 = torch.zeros_like(mom)
    for y0, y1 in [(0, 100), (100, 101), (101, h)]:
        ops.lab3_moments(d[y0:y1].contiguous(), d_pts, d_off, 32, mom2, y0=y0)
    assert torch.equal(mom, mom2)
    params = ops.lab3_finalize(mom)
    want = MahalanobisClassifier.reference_params(img, classes, mode="moments")
    assert np.array_equal(params.cpu().numpy().view(np.int64), want.view(np.int64))   # bitwise


@pytest.mark.parametrize("variant", ["direct", "lut"])
def test_lab3_classify_bit_exact(cuda_device, variant):
    rng = np.random.default_rng(22)
    h, w = 211, 403                                   # odd pixel count: vector body + scalar tail
    img = rng.integers(0, 256, size=(h, w, 4), dtype=np.uint8)
    classes = random_classes(rng, w, h, 32, 300)
    params = MahalanobisClassifier.reference_params(img, classes, mode="moments")
    want = MahalanobisClassifier.reference_classify(img, params)
    d = torch.from_numpy(img).to(cuda_device)
    ops.lab3_classify(d, torch.from_numpy(params).to(cuda_device), variant=variant)
    assert np.array_equal(d.cpu().numpy(), want)
    # PyTorch float64 restatement, ignoring numerical near-ties
    rgb = torch.from_numpy(img[..., :3].astype(np.float64)).to(cuda_device)
    p = torch.from_numpy(params).to(cuda_device)
    diff = rgb[None] - p[:, None, None, :3]
    dist = torch.einsum("chwi,cij,chwj->chw", diff, p[:, 3:].reshape(-1, 3, 3), diff)
    top2 = torch.topk(dist, 2, dim=0, largest=False).values
    clear = (top2[1] - top2[0]) > 1e-9 * top2[1].abs()
    got = d[..., 3].to(torch.int64)
    assert torch.equal(got[clear], dist.argmin(dim=0)[clear]) and clear.float().mean() > 0.99


@pytest.mark.parametrize("kind", ["noise", "separated"])
def test_lab3_two_level_apply_matches_one_level(cuda_device, kind):
    """>= 2^22 pixels + full scratch -> the image pass looks colours up in a shared-memory table
    of 4x4x8 colour cells first and in the 16 MiB table only for cells that straddle a class
    boundary.  Must equal the one-level pass (16 MiB scratch only) and the direct fp64 kernel."""
    rng = np.random.default_rng(23)
    h, w = 2048, 2304 + 1                               # 4.7 M pixels, odd count -> scalar tail too
    img = rng.integers(0, 256, size=(h, w, 4), dtype=np.uint8)
    if kind == "noise":                                 # near-identical classes: ragged boundaries
        classes = random_classes(rng, w, h, 32, 300)
    else:                                               # 4 colour blobs sampled from 4 image quadrants
        centres = np.array([[40, 50, 60], [200, 60, 70], [60, 210, 90], [120, 130, 220]])
        for q, (ys, xs) in enumerate([(slice(0, 1024), slice(0, 1152)), (slice(0, 1024), slice(1152, w)),
                                      (slice(1024, h), slice(0, 1152)), (slice(1024, h), slice(1152, w))]):
            blk = img[ys, xs, :3]
            blk[...] = np.clip(centres[q] + rng.normal(0, 12, blk.shape), 0, 255).astype(np.uint8)
        boxes = [(0, 1024, 0, 1152), (0, 1024, 1152, w), (1024, h, 0, 1152), (1024, h, 1152, w)]
        classes = [np.stack([rng.integers(x0, x1, 400), rng.integers(y0, y1, 400)], axis=1) for y0, y1, x0, x1 in boxes]
        img[::7, ::5, :3] = rng.integers(0, 256, size=img[::7, ::5, :3].shape, dtype=np.uint8)   # every cell gets visitors
    d0 = torch.from_numpy(img).to(cuda_device)
    params = ops.lab3_params_from_points(d0, classes)
    full = int(N.lib().b200_lab3_classify_scratch_bytes(h * w, ops.classify.CLASSIFY_VARIANTS["lut"]))
    coarse_bytes = int(N.lib().b200_lab3_coarse_bytes())
    assert full == (1 << 24) + coarse_bytes
    outs = {}
    for name, variant, nbytes in (("two_level", "lut", full), ("one_level", "lut", 1 << 24), ("direct", "direct", 0)):
        d = d0.clone()
        scratch = torch.full((max(nbytes, 1),), 0xEE, dtype=torch.uint8, device=cuda_device)
        N.check(N.lib().b200_lab3_classify(N.ptr(d), h * w, len(classes), N.ptr(params),
                                           ops.classify.CLASSIFY_VARIANTS[variant], N.ptr(scratch) if nbytes else None,
                                           nbytes, N.current_stream_ptr()), "classify")
        outs[name] = d.cpu().numpy()
        if name == "two_level":
            lut = scratch[: 1 << 24].cpu().numpy()
            cells = scratch[1 << 24:].cpu().numpy()
    assert np.array_equal(outs["two_level"], outs["one_level"])
    assert np.array_equal(outs["two_level"], outs["direct"])
    # the cell table: uniform cells name the class of all their 128 colours, the rest say "mixed"
    mixed = cells == 0xFF
    assert 0 < mixed.sum() < cells.size, "expected both uniform and mixed cells"
    cube = lut.reshape(32, 8, 64, 4, 64, 4)            # B = (bc, b), G = (gc, g), R = (rc, r)
    lo, hi = cube.min(axis=(1, 3, 5)), cube.max(axis=(1, 3, 5))
    want_cells = np.where(lo == hi, lo, 0xFF).astype(np.uint8).reshape(-1)   # index = bc<<12 | gc<<6 | rc
    assert np.array_equal(cells, want_cells)
    if kind == "separated":
        assert mixed.mean() < 0.25


def test_lab3_geom_and_golden(cuda_device, repo_root, tmp_path):
    src = ImgData(str(repo_root / "lab3/data/test_01_lab3.txt"), out_dir=str(tmp_path), materialize=False).rgba
    gold = ImgData(str(repo_root / "lab3/data_out_gt/test_01_lab3.txt"), out_dir=str(tmp_path), materialize=False).rgba
    classes = [np.array([[1, 2], [1, 0], [2, 2], [2, 1]]), np.array([[0, 0], [0, 1], [1, 1], [2, 0]])]
    for stats in ("device", "host"):
        d = torch.from_numpy(src.copy()).to(cuda_device)
        MahalanobisClassifier(cuda_device, stats=stats).run_device(d, classes)
        assert np.array_equal(d.cpu().numpy(), gold), stats
    d = torch.from_numpy(src.copy()).to(cuda_device)
    params = ops.lab3_params_from_points(d, classes)
    ops.lab3_classify_geom(d, params, 256, 256)
    assert np.array_equal(d.cpu().numpy(), gold)


# ------------------------------------------------------------------------ lab5
def make_keys(rng, n, dtype, kind):
    if dtype == np.float32:
        k = (rng.standard_normal(n) * 1e4).astype(np.float32)
        if n >= 4:
            k[:4] = [0.0, -0.0, np.inf, -np.inf]
    elif dtype == np.uint8:
        k = rng.integers(0, 256, n, dtype=np.int64).astype(np.uint8)
    else:
        info = np.iinfo(dtype)
        k = rng.integers(info.min, info.max, n, dtype=np.int64, endpoint=True).astype(dtype)
    if kind == "dups":
        k = k[rng.integers(0, max(1, min(n, 17)), n)]
    elif kind == "sorted":
        k = np.sort(k)
    elif kind == "reversed":
        k = np.sort(k)[::-1].copy()
    elif kind == "equal":
        k[:] = k[0]
    return k


@pytest.mark.parametrize("dtype", [np.int32, np.uint32, np.float32, np.uint8])
def test_sort_matches_torch_and_oracle(cuda_device, dtype):
    rng = np.random.default_rng(31)
    for n in [1, 2, 10, 8191, 8192, 8193, 100_003, 3_000_001]:
        for kind in (["random"] if n > 200_000 else ["random", "dups", "sorted", "reversed", "equal"]):
            k = make_keys(rng, n, dtype, kind)
            d = torch.from_numpy(k.copy()).to(cuda_device)
            ops.sort_(d)
            got = d.cpu().numpy()
            assert np.array_equal(got, np.sort(k)), (dtype, n, kind)
            assert got.tobytes() == RadixSorter.reference(k).tobytes() or dtype == np.float32
            if dtype != np.uint32:
                want_t = torch.sort(torch.from_numpy(k.copy()).to(cuda_device)).values
                assert torch.equal(d, want_t)


def test_sort_runs_of_equal_digits(cuda_device):
    """Warps whose keys share a digit (per half or whole) take the atomic-free ranking path;
    run lengths around 256 / 512 keys put the boundary inside, between and across warps."""
    rng = np.random.default_rng(32)
    n = 2_500_003
    for run in [1, 100, 256, 300, 512, 8192, 100_000, n]:
        hi = np.repeat(rng.integers(0, 1 << 24, n // run + 1, dtype=np.int64), run)[:n]
        for low_bits in (0, 8):
            k = ((hi << 8) | rng.integers(0, 1 << low_bits, n, dtype=np.int64)).astype(np.uint32).view(np.int32)
            d = torch.from_numpy(k.copy()).to(cuda_device)
            ops.sort_(d)
            assert np.array_equal(d.cpu().numpy(), np.sort(k)), (run, low_bits)


def test_sort_passes_with_one_digit_value(cuda_device):
    """A pass whose digit is the same for every key only copies (flag from the scan kernel):
    small-range integers, floats of one exponent, and the neighbours that must NOT take it."""
    rng = np.random.default_rng(33)
    for n in [5000, 1_000_003]:
        cases = [
            rng.integers(0, 1000, n).astype(np.int32),           # passes 2, 3 constant
            rng.integers(-500, 500, n).astype(np.int32),         # top digits 0x00 / 0xff: two bins
            (rng.integers(0, 256, n) << 16).astype(np.int32),    # passes 0, 1, 3 constant
            (rng.integers(0, 2, n) << 31 | 7).astype(np.uint32).view(np.int32),  # two bins in the last pass
            (1.0 + rng.random(n)).astype(np.float32),            # one exponent
            np.full(n, -3, dtype=np.int32),
        ]
        for k in cases:
            d = torch.from_numpy(k.copy()).to(cuda_device)
            ops.sort_(d)
            assert np.array_equal(d.cpu().numpy(), np.sort(k)), (n, k.dtype, k[:4])


def test_sort_fixtures_and_host_pipeline(cuda_device, repo_root):
    for name, dt in [("int10", np.int32), ("float10", np.float32), ("uchar10", np.uint8)]:
        blob = (repo_root / "lab5/data" / name).read_bytes()
        k = np.frombuffer(blob, dtype=dt, offset=4)
        out = RadixSorter(cuda_device).run_host(torch.from_numpy(k.copy()))
        assert np.array_equal(out.numpy(), np.sort(k))


def test_vector_sub_host_pipeline(cuda_device):
    a = torch.randn(100_000, dtype=torch.float64).pin_memory()
    b = torch.randn(100_000, dtype=torch.float64).pin_memory()
    assert torch.equal(VectorSub(cuda_device).run_host(a, b), a - b)
