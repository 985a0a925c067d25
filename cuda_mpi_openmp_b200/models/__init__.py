"""The lab workloads as pipelines: host data in -> sm_100a kernels -> host data out.

These are the public entry points a user (and bench.py's end-to-end arm) calls:

=====================  ======================================================
``VectorSub``          lab1  c = a - b over doubles
``RobertsFilter``      lab2  Roberts-cross edge filter on RGBA8 images
``MahalanobisClassifier``  lab3  class statistics + per-pixel classification
``RadixSorter``        lab5  ascending sort of int32/uint32/float32/uint8 keys
=====================  ======================================================
"""
from .vector_sub import VectorSub
from .roberts import RobertsFilter
from .classifier import MahalanobisClassifier
from .sorter import RadixSorter

__all__ = ["VectorSub", "RobertsFilter", "MahalanobisClassifier", "RadixSorter"]
