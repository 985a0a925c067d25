import os
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def repo_root() -> Path:
    return ROOT


@pytest.fixture(scope="session")
def built(repo_root):
    """Native library + contract binaries, built once per session (nvcc cross-compiles
    without a GPU; on the GPU box the snapshot already carries the artefacts)."""
    res = subprocess.run(["make", "-C", str(repo_root), "-j", str(os.cpu_count() or 4), "all"],
                         capture_output=True, text=True)
    assert res.returncode == 0, res.stdout + res.stderr
    return repo_root


@pytest.fixture(scope="session")
def lib(built):
    from cuda_mpi_openmp_b200 import _native

    return _native.lib()


@pytest.fixture(scope="session")
def cuda_device(lib):
    import torch

    if not torch.cuda.is_available():
        pytest.fail("test marked gpu but no CUDA device is visible")
    return torch.device("cuda:0")
