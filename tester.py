"""Import shim: `from tester import BaseTester, BaseLabProcessor, ...` keeps working
(the implementation lives in cuda_mpi_openmp_b200/harness/core.py)."""
from cuda_mpi_openmp_b200.harness.core import *  # noqa: F401,F403
from cuda_mpi_openmp_b200.harness.core import (BaseLabProcessor, BaseTester, SubProcessResult, TaskResult,  # noqa: F401
                                                TIME_KERNEL_EXE_PATTERN, parse_kernel_time, print_stat_time,
                                                run_subprocess)
