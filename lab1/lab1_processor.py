"""Import shim: the processor lives in cuda_mpi_openmp_b200/harness/processors.py."""
from cuda_mpi_openmp_b200.harness.processors import Lab1Processor  # noqa: F401
