"""Contract checks of the CPU binaries and utilities (stdin/stdout formats)."""
import struct
import subprocess

import numpy as np
import pytest


def run(exe, stdin, binary=False):
    r = subprocess.run([str(exe)], input=stdin, capture_output=True, text=not binary, timeout=60)
    assert r.returncode == 0, r.stderr
    return r.stdout


@pytest.mark.parametrize("stdin,want", [
    ("1 -3 2", "2.000000 1.000000\n"), ("0 0 0", "any\n"), ("0 0 5", "incorrect\n"),
    ("0 2 1", "-0.500000\n"), ("1 2 1", "-1.000000\n"), ("1 0 1", "imaginary\n"),
])
def test_hw1(built, stdin, want):
    assert run(built / "hw1/src/hw1", stdin) == want


def test_hw2(built):
    assert run(built / "hw2/src/hw2", "5\n3 1 2 5 4") == "1.000000e+00 2.000000e+00 3.000000e+00 4.000000e+00 5.000000e+00 \n"
    assert run(built / "hw2/src/hw2", "0\n") == "\n"


def test_lab1_cpu_contract(built):
    out = run(built / "lab1/src/cpu_exe_to_plot", "3\n1 2 3.5\n0.5 1e100 -2")
    first, rest = out.split("\n", 1)
    assert first.startswith("CPU execution time: <") and first.endswith(" ms>")
    assert rest == "5.0000000000e-01 -1.0000000000e+100 5.5000000000e+00 "
    assert run(built / "lab1/src/cpu_exe", "2\n1 2\n3 5") == "-2.0000000000e+00 -3.0000000000e+00 "


def test_lab3_read_input(built):
    out = run(built / "lab3/src/test_read_input", "2\n2 1 2 3 4\n1 9 8")
    assert "class 1: 2 points" in out and "(3, 4)" in out and "class 2: 1 points" in out and "(9, 8)" in out


@pytest.mark.parametrize("name,exe,fmt", [("int10", "int", "<10i"), ("float10", "float", "<10f"), ("uchar10", "uchar", "<10B")])
def test_lab5_cpu_fixture(built, name, exe, fmt):
    blob = (built / "lab5/data" / name).read_bytes()
    out = run(built / f"lab5/src/cpu_exe_{exe}", blob, binary=True)
    vals = struct.unpack(fmt, out)
    keys = struct.unpack(fmt, blob[4:])
    assert list(vals) == sorted(keys)


def test_lab2_cpu_error_on_missing_file(built):
    r = subprocess.run([str(built / "lab2/src/cpu_exe")], input="/nonexistent.data\n/tmp/x.data", capture_output=True, text=True)
    assert r.returncode != 0 and "cannot open" in r.stderr


def test_threaded_cpu_binaries_match_the_serial_library_oracles(built, lib, tmp_path):
    """The stand-alone CPU binaries split their loops over std::thread workers above a size
    threshold (csrc/oracle.cpp, -DB200_ORACLE_THREADS); libb200labs carries the serial build.
    Inputs big enough to be split must give the same bytes either way."""
    from cuda_mpi_openmp_b200 import _native as N

    rng = np.random.default_rng(77)
    # lab2: 301 x 431 px (> 2^16 pixels -> several row bands)
    h, w = 301, 431
    img = rng.integers(0, 256, size=(h, w, 4), dtype=np.uint8)
    src, dst = tmp_path / "in.data", tmp_path / "out.data"
    src.write_bytes(struct.pack("<ii", w, h) + img.tobytes())
    run(built / "lab2/src/cpu_exe", f"{src}\n{dst}\n")
    want = np.empty_like(img)
    lib.b200_cpu_roberts(N.ptr(img), N.ptr(want), w, h)
    got = np.frombuffer(dst.read_bytes()[8:], dtype=np.uint8).reshape(h, w, 4)
    assert np.array_equal(got, want)
    # lab3: same image, two classes of 40 points each, class byte in alpha
    pts = [rng.integers(0, [w, h], size=(40, 2)) for _ in range(2)]
    spec = "\n".join(f"{len(p)} " + " ".join(f"{x} {y}" for x, y in p) for p in pts)
    dst3 = tmp_path / "out3.data"
    run(built / "lab3/src/cpu_exe", f"{src}\n{dst3}\n2\n{spec}\n")
    got3 = np.frombuffer(dst3.read_bytes()[8:], dtype=np.uint8).reshape(h, w, 4)
    from cuda_mpi_openmp_b200.models import MahalanobisClassifier

    params = MahalanobisClassifier.reference_params(img, pts, mode="host")
    assert np.array_equal(got3, MahalanobisClassifier.reference_classify(img, params))
    # lab1: 200k elements (> 2^16 -> several slices)
    n = 200_000
    a, b = rng.uniform(-1e100, 1e100, n), rng.uniform(-1e100, 1e100, n)
    text = f"{n}\n" + " ".join(f"{x:.10e}" for x in a) + "\n" + " ".join(f"{x:.10e}" for x in b) + "\n"
    out = run(built / "lab1/src/cpu_exe", text)
    a_p = np.array([float(f"{x:.10e}") for x in a])
    b_p = np.array([float(f"{x:.10e}") for x in b])
    assert out.split() == [f"{x:.10e}" for x in (a_p - b_p)]
