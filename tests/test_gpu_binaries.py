"""The CUDA contract binaries and the harness driving them, on a real GPU."""
import os
import struct
import subprocess
import sys

import numpy as np
import pandas as pd
import pytest

from cuda_mpi_openmp_b200.models import RobertsFilter
from cuda_mpi_openmp_b200.utils import ImgData, encode_data_bytes

pytestmark = pytest.mark.gpu


def run(exe, stdin, binary=False, env=None, args=()):
    r = subprocess.run([str(exe), *args], input=stdin, capture_output=True, text=not binary, timeout=120,
                       env={**os.environ, **(env or {})})
    return r


def test_gpu_info_labels(built, cuda_device):
    out = run(built / "gpu_info/src/gpu_info", "").stdout.splitlines()
    labels = [l.split(" : ")[0] for l in out]
    assert labels == ["Compute capability", "Name", "Total Global Memory", "Shared memory per block",
                      "Max threads per block", "Max block", "Total constant memory", "Multiprocessors count"]
    ext = run(built / "gpu_info/src/gpu_info", "", args=("--extended",)).stdout
    assert ext.splitlines()[:8] == out and "L2 cache size" in ext


def test_lab1_cuda_binaries(built, cuda_device):
    body = "3\n1 2 3.5\n0.5 1e100 -2"
    want = "5.0000000000e-01 -1.0000000000e+100 5.5000000000e+00 "
    r = run(built / "lab1/src/cuda_exe", body)
    assert r.returncode == 0 and r.stdout == want, r.stderr
    for dims in ("512\n512\n", "1\n32\n", "-1\n-1\n"):
        r = run(built / "lab1/src/cuda_exe_to_plot", dims + body)
        first, rest = r.stdout.split("\n", 1)
        assert first.startswith("CUDA execution time: <") and rest == want
    bad = run(built / "lab1/src/cuda_exe_to_plot", "1\n4096\n" + body)
    assert bad.returncode == 1 and "[ERROR CUDA]" in bad.stderr


@pytest.mark.parametrize("env", [{}, {"B200LABS_ROBERTS": "generic"}, {"B200LABS_WARM": "1"}])
def test_lab2_cuda_binaries(built, cuda_device, tmp_path, env):
    for name in ("test_01", "test_02"):
        src = ImgData(str(built / "lab2/data" / f"{name}.txt"), out_dir=str(tmp_path))
        gold = ImgData(str(built / "lab2/data_out_gt" / f"{name}.txt"), out_dir=str(tmp_path / "gt"))
        out = tmp_path / f"{name}_out.data"
        r = run(built / "lab2/src/cuda_exe", f"{src.c_data_bytes_path}\n{out}", env=env)
        assert r.returncode == 0 and r.stdout == "" and out.read_bytes() == gold.c_data_bytes
        r = run(built / "lab2/src/to_plot_cuda_exe", f"32\n32\n16\n16\n{src.c_data_bytes_path}\n{out}", env=env)
        assert r.stdout.startswith("CUDA execution time: <") and r.stdout.endswith("FINISHED!\n")
        assert out.read_bytes() == gold.c_data_bytes
    big = np.random.default_rng(0).integers(0, 256, size=(600, 1024, 4), dtype=np.uint8)
    p = tmp_path / "big.data"
    p.write_bytes(encode_data_bytes(big))
    for tool, dims in (("cuda_exe", ""), ("to_plot_cuda_exe", "-1\n-1\n-1\n-1\n"), ("to_plot_cuda_exe", "16\n16\n32\n32\n")):
        r = run(built / "lab2/src" / tool, f"{dims}{p}\n{tmp_path / 'big_out.data'}", env=env)
        assert r.returncode == 0, r.stderr
        got = ImgData(str(tmp_path / "big_out.data"), materialize=False).rgba
        assert np.array_equal(got, RobertsFilter.reference(big))


def test_lab3_cuda_binaries(built, cuda_device, tmp_path):
    src = ImgData(str(built / "lab3/data/test_01_lab3.txt"), out_dir=str(tmp_path))
    gold = ImgData(str(built / "lab3/data_out_gt/test_01_lab3.txt"), out_dir=str(tmp_path / "gt"))
    spec = "2\n4 1 2 1 0 2 2 2 1\n4 0 0 0 1 1 1 2 0"
    out = tmp_path / "l3.data"
    for env in ({}, {"B200LABS_LAB3_STATS": "host"}):
        r = run(built / "lab3/src/cuda_exe", f"{src.c_data_bytes_path}\n{out}\n{spec}", env=env)
        assert r.returncode == 0 and r.stdout == "" and out.read_bytes() == gold.c_data_bytes, r.stderr
    r = run(built / "lab3/src/to_plot_cuda_exe", f"256\n256\n{src.c_data_bytes_path}\n{out}\n{spec}")
    assert r.stdout.startswith("CUDA execution time: <") and out.read_bytes() == gold.c_data_bytes


def test_lab5_cuda_binaries(built, cuda_device, tmp_path):
    for name, exe, fmt in [("int10", "int", "<10i"), ("float10", "float", "<10f"), ("uchar10", "uchar", "<10B")]:
        blob = (built / "lab5/data" / name).read_bytes()
        r = run(built / f"lab5/src/cuda_exe_{exe}", blob, binary=True)
        assert r.returncode == 0, r.stderr
        assert list(struct.unpack(fmt, r.stdout)) == sorted(struct.unpack(fmt, blob[4:]))
    keys = np.random.default_rng(1).integers(-2**31, 2**31 - 1, 200_000, dtype=np.int64).astype(np.int32)
    p = tmp_path / "int_big"
    p.write_bytes(struct.pack("<i", keys.size) + keys.tobytes())
    r = run(built / "lab5/src/to_plot_cuda_exe", f"512\n512\n{p}\n{tmp_path / 'sorted'}\nint")
    assert r.stdout.startswith("CUDA execution time: <"), r.stderr
    got = np.frombuffer((tmp_path / "sorted").read_bytes(), dtype=np.int32, offset=4)
    assert np.array_equal(got, np.sort(keys))


def test_large_files_take_the_pipelined_path(built, cuda_device, tmp_path):
    """Payloads of 32 MiB and more are streamed file -> pinned ring -> GPU -> pinned ring -> file by
    several I/O threads (apps/cli_pipeline.h); the bytes must be the ones the plain path writes."""
    rng = np.random.default_rng(9)
    img = rng.integers(0, 256, size=(2900, 3004, 4), dtype=np.uint8)         # 34.8 MB, odd chunk tail
    p, out = tmp_path / "big.data", tmp_path / "big_out.data"
    p.write_bytes(encode_data_bytes(img))
    r = run(built / "lab2/src/to_plot_cuda_exe", f"-1\n-1\n-1\n-1\n{p}\n{out}")
    assert r.returncode == 0 and r.stdout.endswith("FINISHED!\n"), r.stderr
    assert np.array_equal(ImgData(str(out), materialize=False).rgba, RobertsFilter.reference(img))
    # in == out forces the plain path (reading must finish before the file is truncated)
    r = run(built / "lab2/src/cuda_exe", f"{p}\n{p}")
    assert r.returncode == 0 and p.read_bytes() == out.read_bytes()
    # lab3: same image, classes from random points; piped output == host-statistics-free plain path
    p.write_bytes(encode_data_bytes(img))
    pts = "\n".join("50 " + " ".join(f"{x} {y}" for x, y in zip(rng.integers(0, 3004, 50), rng.integers(0, 2900, 50)))
                    for _ in range(5))
    r = run(built / "lab3/src/cuda_exe", f"{p}\n{out}\n5\n{pts}")
    assert r.returncode == 0, r.stderr
    got = ImgData(str(out), materialize=False).rgba
    small = tmp_path / "same_in_out.data"
    small.write_bytes(encode_data_bytes(img))
    r = run(built / "lab3/src/cuda_exe", f"{small}\n{small}\n5\n{pts}")          # in == out: plain path
    assert r.returncode == 0, r.stderr
    assert np.array_equal(got, ImgData(str(small), materialize=False).rgba)
    assert np.array_equal(got[..., :3], img[..., :3]) and got[..., 3].max() < 5
    # lab5: 9 M keys = 36 MB
    keys = rng.integers(-2**31, 2**31 - 1, 9_000_001, dtype=np.int64).astype(np.int32)
    kp = tmp_path / "int_huge"
    kp.write_bytes(struct.pack("<i", keys.size) + keys.tobytes())
    r = run(built / "lab5/src/to_plot_cuda_exe", f"1\n1\n{kp}\n{tmp_path / 'sorted_huge'}\nint")
    assert r.returncode == 0 and r.stdout.startswith("CUDA execution time: <"), r.stderr
    raw = (tmp_path / "sorted_huge").read_bytes()
    assert struct.unpack("<i", raw[:4])[0] == keys.size and len(raw) == 4 + 4 * keys.size
    assert np.array_equal(np.frombuffer(raw, dtype=np.int32, offset=4), np.sort(keys))


@pytest.mark.parametrize("lab,exe,sizes,extra", [
    ("lab1", "cuda_exe_to_plot", "[[1, 32], [512, 512], [1024, 1024]]", ["--binary_path_cpu", "./lab1/src/cpu_exe_to_plot"]),
    ("lab2", "to_plot_cuda_exe", "[[[32, 32], [16, 16]], [[16, 16], [32, 32]]]", ["--binary_path_cpu", "./lab2/src/cpu_exe"]),
    ("lab3", "to_plot_cuda_exe", "[[256, 256]]", []),
    ("lab5", "to_plot_cuda_exe", "[[512, 512]]", ["--extra_random", "1"]),
])
def test_harness_drives_cuda_binaries(built, cuda_device, lab, exe, sizes, extra):
    cmd = [sys.executable, "run_test.py", "--binary_path_cuda", f"./{lab}/src/{exe}", "--k_times", "7",
           "--kernel_sizes", sizes, "--metadata_columns2plot", "[]"] + extra
    r = subprocess.run(cmd, cwd=built, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "SUCCESS!" in r.stdout and "FAILED" not in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
    stats = pd.read_csv(built / lab / "src" / f"stats_{exe}.csv")
    assert stats["test_verification_result"].all() and stats["time_kernel_exe_ms"].notna().all()
