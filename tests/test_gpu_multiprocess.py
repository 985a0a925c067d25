"""One process per GPU over NCCL/NVLink: fused peer-memory kernels vs the CPU oracle
and vs the NCCL baseline.  Needs >= 2 GPUs: skipped on a single-GPU box, unless
B200LABS_REQUIRE_MGPU=1 says the box is supposed to have them (then it fails loudly)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu


def test_sharded_labs_over_all_visible_gpus(built, cuda_device):
    n = torch.cuda.device_count()
    if n < 2:
        if os.environ.get("B200LABS_REQUIRE_MGPU") == "1":
            pytest.fail(f"B200LABS_REQUIRE_MGPU=1 but only {n} GPU is visible")
        pytest.skip("needs at least 2 GPUs")
    world = 8 if n >= 8 else 4 if n >= 4 else 2
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", "29533", str(built / "tests/mp_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=built,
                       env={**os.environ, "NCCL_DEBUG": "WARN"})
    log_dir = built / "gpurun_out"                  # full worker output survives the GPU call
    log_dir.mkdir(exist_ok=True)
    (log_dir / "mp_worker_stdout.log").write_text(r.stdout)
    (log_dir / "mp_worker_stderr.log").write_text(r.stderr)
    assert r.returncode == 0 and "MP_WORKER_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
