"""lab1: element-wise double subtraction (ref: lab1/src/main.cu:22-29)."""
from __future__ import annotations

import torch

from .. import _native as N


def _check(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor | None):
    if a.dtype != torch.float64 or b.dtype != torch.float64:
        raise TypeError("vecsub works on float64 tensors (the lab's `double` vectors)")
    if a.shape != b.shape or a.dim() != 1:
        raise ValueError("vecsub needs two 1-D tensors of equal length")
    if not a.is_cuda or not b.is_cuda:
        raise ValueError("vecsub needs CUDA tensors")
    if out is None:
        out = torch.empty_like(a)
    return out


def vecsub(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """c = a - b with the tuned B200 kernel (256-bit streaming accesses)."""
    out = _check(a, b, out)
    N.check(N.lib().b200_vecsub_f64(N.ptr(a), N.ptr(b), N.ptr(out), a.numel(), N.current_stream_ptr()),
            "b200_vecsub_f64")
    return out


def vecsub_geom(a, b, grid: int, block: int, out=None) -> torch.Tensor:
    """c = a - b launched as <<<grid, block>>> (ref: lab1/src/to_plot.cu:72)."""
    out = _check(a, b, out)
    N.check(N.lib().b200_vecsub_f64_geom(N.ptr(a), N.ptr(b), N.ptr(out), a.numel(), grid, block,
                                         N.current_stream_ptr()), "b200_vecsub_f64_geom")
    return out
