"""Black-box experiment runner: launches the contract binaries as subprocesses,
scrapes the ``execution time: <X ms>`` line, verifies outputs against golden data
and writes per-binary CSVs plus a median bar chart.  Same command line and result
files as the reference's run_test.py / tester.py; see harness/core.py."""
from .core import (
    TIME_KERNEL_EXE_PATTERN, parse_kernel_time, TaskResult, SubProcessResult, BaseLabProcessor,
    BaseTester, run_subprocess,
)
from .argkw import unknown_args_to_kwargs
from .processors import Lab1Processor, Lab2Processor, Lab3Processor, Lab5Processor, LAB_PROCESSORS

__all__ = [
    "TIME_KERNEL_EXE_PATTERN", "parse_kernel_time", "TaskResult", "SubProcessResult",
    "BaseLabProcessor", "BaseTester", "run_subprocess", "unknown_args_to_kwargs",
    "Lab1Processor", "Lab2Processor", "Lab3Processor", "Lab5Processor", "LAB_PROCESSORS",
]
