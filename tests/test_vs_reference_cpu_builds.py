"""CPU utilities against the unmodified reference programs built with gcc
(baseline/_ref/bin: hw1, hw2, lab1_cpu_exe, lab2_cpu_exe).  Skipped without them."""
import subprocess

import numpy as np
import pytest

from cuda_mpi_openmp_b200.utils import ImgData


def ref_bin(root, name):
    p = root / "baseline/_ref/bin" / name
    if not p.exists():
        pytest.skip("reference build not present (run baseline/install_reference.sh)")
    return p


def run(exe, stdin):
    r = subprocess.run([str(exe)], input=stdin, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    return r.stdout


@pytest.mark.parametrize("stdin", ["1 -3 2", "0 0 0", "0 0 7", "0 4 2", "1 2 1", "2 0 8", "1e-3 5 1e-3",
                                   "3.5 -2.25 -7.125", "1 -2000 1", "-1 0 4"])
def test_hw1_same_stdout(built, stdin):
    assert run(built / "hw1/src/hw1", stdin) == run(ref_bin(built, "hw1"), stdin)


def test_hw2_same_stdout(built):
    rng = np.random.default_rng(0)
    for n in (0, 1, 2, 17, 300):
        vals = rng.standard_normal(n) * 10.0 ** int(rng.integers(-3, 6))
        stdin = f"{n}\n" + " ".join(f"{v:.7g}" for v in vals)
        assert run(built / "hw2/src/hw2", stdin) == run(ref_bin(built, "hw2"), stdin)


def test_lab1_cpu_same_vector(built):
    rng = np.random.default_rng(1)
    n = 2000
    a, b = rng.uniform(-1e100, 1e100, n), rng.uniform(-1e100, 1e100, n)
    stdin = f"{n}\n" + " ".join(f"{v:.10e}" for v in a) + "\n" + " ".join(f"{v:.10e}" for v in b)
    ours = run(built / "lab1/src/cpu_exe_to_plot", stdin).split("\n", 1)
    ref = run(ref_bin(built, "lab1_cpu_exe"), stdin).split("\n", 1)
    assert ours[0].startswith("CPU execution time: <") and ref[0].startswith("CPU execution time: <")
    assert ours[1] == ref[1]


def test_lab2_cpu_goldens_and_lenna(built, tmp_path):
    ref = ref_bin(built, "lab2_cpu_exe")
    for path in ("lab2/data/test_01.txt", "lab2/data/test_02.txt", "lab2/test_data/lenna.data"):
        src = ImgData(str(built / path), out_dir=str(tmp_path), materialize=False)
        a, b = tmp_path / "ref.data", tmp_path / "ours.data"
        run(ref, f"{src.c_data_bytes_path}\n{a}")
        run(built / "lab2/src/cpu_exe", f"{src.c_data_bytes_path}\n{b}")
        # the reference's gcc build is not FMA-contracted; on these inputs that makes no difference
        assert a.read_bytes() == b.read_bytes(), path
