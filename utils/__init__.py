"""Import shim: `from utils import ImgData, download_file` (see cuda_mpi_openmp_b200/utils)."""
from cuda_mpi_openmp_b200.utils import ImgData, download_file  # noqa: F401
