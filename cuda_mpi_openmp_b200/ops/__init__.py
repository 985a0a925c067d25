"""Kernel-level operations on CUDA tensors (hand-written sm_100a kernels)."""
from .vecsub import vecsub, vecsub_geom
from .roberts import roberts, roberts_geom, ROBERTS_VARIANTS
from .classify import (
    lab3_moments, lab3_finalize, lab3_classify, lab3_classify_geom, lab3_params_from_points,
    CLASSIFY_VARIANTS,
)
from .sort import sort_, sort_scratch_bytes, KEY_TYPES

__all__ = [
    "vecsub", "vecsub_geom", "roberts", "roberts_geom", "ROBERTS_VARIANTS",
    "lab3_moments", "lab3_finalize", "lab3_classify", "lab3_classify_geom",
    "lab3_params_from_points", "CLASSIFY_VARIANTS", "sort_", "sort_scratch_bytes", "KEY_TYPES",
]
