"""Import shim (ref: lab3/img_data_classifier.py)."""
from cuda_mpi_openmp_b200.harness.processors import (GroundTruthClass, MAX_CLASSES, MAX_NUM_POINTS,  # noqa: F401
                                                      get_random_pts)
