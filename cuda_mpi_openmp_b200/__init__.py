"""B200-native GPU lab suite with the capabilities of KoryakovDmitry/cuda-mpi-openmp.

Layout
------
ops/       thin, typed wrappers over the hand-written sm_100a kernels in libb200labs
models/    the lab workloads as pipelines (host data in -> kernels -> host data out)
parallel/  peer-memory runtime and the fused multi-GPU variants (+ NCCL baselines)
utils/     image-format bridge, timing, clock sampling, downloads
harness/   the experiment runner / verifier (run_test.py front end)
"""
__version__ = "0.1.0"

from . import _native  # noqa: F401
