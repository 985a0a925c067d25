"""bench.py's driver contract, as far as it can be checked without a GPU."""
import json
import subprocess
import sys

import bench


def test_defaults_finish_quickly_and_name_the_baseline_config():
    sys.argv = ["bench.py"]
    a = bench.parse_args()
    assert (a.gpus, a.impl, a.size) == (1, "own", 65536) and a.warmup >= 3 and a.steps >= 1
    assert bench.METRIC == "lab2_roberts_effective_GBps" and bench.BYTES_PER_PX == 8


def test_reference_arm_reports_unavailable_without_a_gpu(built):
    r = subprocess.run([sys.executable, "bench.py", "--impl", "reference", "--size", "1024", "--steps", "1",
                        "--warmup", "1"], cwd=built, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-1000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference"
    # on a GPU-less box the unmodified binary fails at its first CUDA call (or is not installed)
    assert "unavailable" in line or line.get("metric") == bench.METRIC
