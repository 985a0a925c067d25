"""CPU oracles against the reference's golden files and independent numpy restatements."""
import numpy as np
import pytest

from cuda_mpi_openmp_b200.models import MahalanobisClassifier, RadixSorter, RobertsFilter, VectorSub
from cuda_mpi_openmp_b200.utils import ImgData

DEFAULT_CLASSES = [np.array([[1, 2], [1, 0], [2, 2], [2, 1]]), np.array([[0, 0], [0, 1], [1, 1], [2, 0]])]


def load(path, tmp_path):
    return ImgData(str(path), out_dir=str(tmp_path), materialize=False).rgba


@pytest.mark.parametrize("name", ["test_01", "test_02"])
def test_roberts_golden(lib, repo_root, tmp_path, name):
    src = load(repo_root / "lab2/data" / f"{name}.txt", tmp_path)
    gold = load(repo_root / "lab2/data_out_gt" / f"{name}.txt", tmp_path)
    assert np.array_equal(RobertsFilter.reference(src), gold)


def roberts_numpy(rgba):
    """float64 restatement (every fp32 step re-rounded); within 1 LSB of the oracle."""
    f = rgba.astype(np.float64)
    c299, c587, c114 = (np.float64(np.float32(v)) for v in (0.299, 0.587, 0.114))
    t = (c587 * f[..., 1]).astype(np.float32).astype(np.float64)
    t = (c299 * f[..., 0] + t).astype(np.float32).astype(np.float64)
    y = (c114 * f[..., 2] + t).astype(np.float32).astype(np.float64)
    yp = np.pad(y, ((0, 1), (0, 1)), mode="edge")
    gx = (yp[1:, 1:] - yp[:-1, :-1]).astype(np.float32).astype(np.float64)
    gy = (yp[:-1, 1:] - yp[1:, :-1]).astype(np.float32).astype(np.float64)
    s = (gy * gy + (gx * gx).astype(np.float32).astype(np.float64)).astype(np.float32)
    g = np.minimum(np.sqrt(s.astype(np.float32)), np.float32(255.0)).astype(np.uint8)
    out = np.repeat(g[..., None], 4, axis=2)
    out[..., 3] = rgba[..., 3]
    return out


def test_roberts_oracle_vs_numpy(lib, repo_root, tmp_path):
    rng = np.random.default_rng(0)
    for shape in [(1, 1), (1, 7), (5, 1), (37, 53), (128, 260)]:
        img = rng.integers(0, 256, size=shape + (4,), dtype=np.uint8)
        a, b = RobertsFilter.reference(img), roberts_numpy(img)
        assert np.array_equal(a[..., 3], img[..., 3])
        assert np.abs(a.astype(int) - b.astype(int)).max() <= 1
    real = load(repo_root / "lab2/test_data/lenna.data", tmp_path)
    a, b = RobertsFilter.reference(real), roberts_numpy(real)
    assert (a != b).mean() < 1e-3


def test_lab3_golden_both_stat_modes(lib, repo_root, tmp_path):
    src = load(repo_root / "lab3/data/test_01_lab3.txt", tmp_path)
    gold = load(repo_root / "lab3/data_out_gt/test_01_lab3.txt", tmp_path)
    for mode in ("moments", "ref"):
        params = MahalanobisClassifier.reference_params(src, DEFAULT_CLASSES, mode=mode)
        assert np.array_equal(MahalanobisClassifier.reference_classify(src, params), gold), mode


def lab3_numpy(rgba, classes):
    rgb = rgba[..., :3].astype(np.float64)
    dists = []
    for pts in classes:
        s = np.stack([rgb[y, x] for x, y in pts])
        mu = s.mean(axis=0)
        cov = (s - mu).T @ (s - mu) / (len(pts) - 1)
        inv = np.linalg.inv(cov)
        d = rgb - mu
        dists.append(np.einsum("hwi,ij,hwj->hw", d, inv, d))
    return np.argmin(np.stack(dists), axis=0), np.stack(dists)


def test_lab3_moments_vs_two_pass_and_numpy(lib):
    rng = np.random.default_rng(1)
    img = rng.integers(0, 256, size=(64, 48, 4), dtype=np.uint8)
    classes = [np.stack([rng.integers(0, 48, 200), rng.integers(0, 64, 200)], axis=1) for _ in range(5)]
    p_m = MahalanobisClassifier.reference_params(img, classes, mode="moments")
    p_r = MahalanobisClassifier.reference_params(img, classes, mode="ref")
    assert np.array_equal(p_m[:, :3], p_r[:, :3])          # means are bit-identical
    assert np.allclose(p_m, p_r, rtol=1e-10, atol=0)        # inverses agree to rounding
    got = MahalanobisClassifier.reference_classify(img, p_m)[..., 3]
    want, d = lab3_numpy(img, classes)
    part = np.partition(d, 1, axis=0)
    clear = (part[1] - part[0]) > 1e-9 * np.abs(part[1])   # ignore numerical near-ties
    assert np.array_equal(got[clear], want[clear]) and clear.mean() > 0.99


def test_lab3_degenerate_single_point(lib):
    img = np.full((2, 2, 4), 7, dtype=np.uint8)
    params = MahalanobisClassifier.reference_params(img, [np.array([[0, 0]])], mode="moments")
    out = MahalanobisClassifier.reference_classify(img, params)
    assert (out[..., 3] == 255).all()   # NaN distances never win: class -1 -> 0xFF, like the reference


def test_sort_oracle(lib):
    rng = np.random.default_rng(2)
    for dt in (np.int32, np.uint32, np.uint8):
        info = np.iinfo(dt)
        k = rng.integers(info.min, info.max, size=1000, dtype=np.int64, endpoint=True).astype(dt)
        assert np.array_equal(RadixSorter.reference(k), np.sort(k))
    f = (rng.standard_normal(1000) * 1e3).astype(np.float32)
    f[:4] = [0.0, -0.0, np.inf, -np.inf]
    assert np.array_equal(RadixSorter.reference(f), np.sort(f))


def test_vecsub_reference():
    a, b = np.array([1.0, 2.0, 1e100]), np.array([0.5, 1e100, -1e100])
    assert np.array_equal(VectorSub.reference(a, b), a - b)
