"""Import shim for the reference's module name (sic: `hundle_unkown`)."""
from cuda_mpi_openmp_b200.harness.argkw import unknown_args_to_kwargs

hundle_unkown = unknown_args_to_kwargs
