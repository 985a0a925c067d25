"""ctypes binding of libb200labs.so (C ABI declared in include/b200labs.h).

The library is built in-tree by ``make lib`` (or ``__graft_entry__.build()``) and
lives at ``cuda_mpi_openmp_b200/lib/libb200labs.so``.  Loading is lazy; on a box
with a GPU a missing library is a hard error (no silent PyTorch fallback).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import threading
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
REPO_ROOT = PKG_DIR.parent
# B200LABS_LIB points at another build of the library (kernel-variant experiments)
LIB_PATH = Path(os.environ["B200LABS_LIB"]) if os.environ.get("B200LABS_LIB") else PKG_DIR / "lib" / "libb200labs.so"

_lock = threading.Lock()
_lib = None

c_void_p, c_int, c_int64, c_size_t = C.c_void_p, C.c_int, C.c_int64, C.c_size_t

# name -> (restype, argtypes)
_SIGNATURES = {
    "b200_last_error": (C.c_char_p, []),
    "b200_version": (c_int, []),
    "b200_launch_count": (C.c_uint64, []),
    "b200_gpu_info": (c_int, [c_int, C.c_char_p, c_int, c_int]),
    "b200_flush_l2": (c_int, [c_void_p, c_size_t, c_void_p]),
    "b200_vecsub_f64": (c_int, [c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "b200_vecsub_f64_geom": (c_int, [c_void_p, c_void_p, c_void_p, c_size_t, c_int, c_int, c_void_p]),
    "b200_roberts": (c_int, [c_void_p, c_void_p, c_int, c_int64, c_void_p, c_int, c_void_p]),
    "b200_roberts_geom": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "b200_lab3_moments": (c_int, [c_void_p, c_int, c_int64, c_int64, c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "b200_lab3_finalize": (c_int, [c_void_p, c_int, c_void_p, c_void_p]),
    "b200_lab3_classify_scratch_bytes": (c_size_t, [c_int64, c_int]),
    "b200_lab3_classify": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_int, c_void_p, c_size_t, c_void_p]),
    "b200_lab3_classify_geom": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_int, c_int, c_void_p]),
    "b200_sort_scratch_bytes": (c_size_t, [c_size_t, c_int]),
    "b200_sort": (c_int, [c_void_p, c_void_p, c_size_t, c_int, c_void_p, c_size_t, c_void_p]),
    # CPU oracles (host pointers)
    "b200_cpu_vecsub_f64": (None, [c_void_p, c_void_p, c_void_p, c_size_t]),
    "b200_cpu_roberts": (None, [c_void_p, c_void_p, c_int, c_int]),
    "b200_cpu_lab3_stats_ref": (None, [c_void_p, c_int, c_void_p, c_void_p, c_int, c_void_p]),
    "b200_cpu_lab3_moments": (None, [c_void_p, c_int, c_int64, c_int64, c_void_p, c_void_p, c_int, c_void_p]),
    "b200_cpu_lab3_finalize": (None, [c_void_p, c_int, c_void_p]),
    "b200_cpu_lab3_classify": (None, [c_void_p, c_int64, c_int, c_void_p]),
    "b200_cpu_sort": (None, [c_void_p, c_size_t, c_int]),
}

# optional symbols (multi-GPU runtime etc.) are registered by the modules that use them
_EXTRA_SIGNATURES: dict = {}


class NativeError(RuntimeError):
    pass


def register(name: str, restype, argtypes) -> None:
    _EXTRA_SIGNATURES[name] = (restype, argtypes)
    if _lib is not None and hasattr(_lib, name):
        fn = getattr(_lib, name)
        fn.restype, fn.argtypes = restype, argtypes


def build(verbose: bool = False) -> Path:
    """Compile the native library (and the contract binaries) with nvcc for sm_100a."""
    cmd = ["make", "-C", str(REPO_ROOT), "-j", str(os.cpu_count() or 4), "all"]
    res = subprocess.run(cmd, capture_output=not verbose, text=True)
    if res.returncode != 0:
        raise NativeError(f"native build failed:\n{res.stdout}\n{res.stderr}")
    return LIB_PATH


def lib() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not LIB_PATH.exists():
            if os.environ.get("B200LABS_NO_AUTOBUILD") == "1":
                raise NativeError(f"{LIB_PATH} is missing; run `make lib`")
            build()
        handle = C.CDLL(str(LIB_PATH))
        for name, (res, args) in {**_SIGNATURES, **_EXTRA_SIGNATURES}.items():
            if hasattr(handle, name):
                fn = getattr(handle, name)
                fn.restype, fn.argtypes = res, args
        _lib = handle
    return _lib


def check(code: int, what: str = "") -> None:
    if code != 0:
        msg = lib().b200_last_error().decode(errors="replace")
        raise NativeError(f"{what or 'libb200labs'} failed ({code}): {msg}")


def launch_count() -> int:
    return int(lib().b200_launch_count())


def current_stream_ptr():
    import torch

    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t) -> C.c_void_p:
    """Device (or host) address of a contiguous torch tensor / numpy array."""
    if t is None:
        return C.c_void_p(0)
    if hasattr(t, "data_ptr"):
        if not t.is_contiguous():
            raise ValueError("tensor must be contiguous")
        return C.c_void_p(t.data_ptr())
    if hasattr(t, "ctypes"):
        if not t.flags["C_CONTIGUOUS"]:
            raise ValueError("array must be C-contiguous")
        return C.c_void_p(t.ctypes.data)
    if isinstance(t, int):
        return C.c_void_p(t)
    raise TypeError(f"cannot take the address of {type(t)}")
