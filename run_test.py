#!/usr/bin/env python
"""Harness entry point with the reference's command line (ref: run_test.py).
Run from the repository root:  python run_test.py --binary_path_cuda ./lab2/src/to_plot_cuda_exe ..."""
from cuda_mpi_openmp_b200.harness.cli import main

if __name__ == "__main__":
    main()
