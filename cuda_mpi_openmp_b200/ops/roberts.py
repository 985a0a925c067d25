"""lab2: Roberts-cross edge filter on RGBA8 images (ref: lab2/src/main.cu:15-52)."""
from __future__ import annotations

import torch

from .. import _native as N

ROBERTS_VARIANTS = {"auto": 0, "stream": 1, "tma": 2, "generic": 3, "cluster": 4}


def _as_image(t: torch.Tensor) -> tuple[int, int]:
    """Accepts [h, w, 4] uint8 or [h, w] int32/uint32-like; returns (h, w)."""
    if not t.is_cuda or not t.is_contiguous():
        raise ValueError("roberts needs a contiguous CUDA tensor")
    if t.dtype == torch.uint8 and t.dim() == 3 and t.shape[2] == 4:
        return t.shape[0], t.shape[1]
    if t.dtype in (torch.int32, torch.uint32) and t.dim() == 2:
        return t.shape[0], t.shape[1]
    raise TypeError("image must be uint8 [h, w, 4] or int32 [h, w]")


def roberts(img: torch.Tensor, out: torch.Tensor | None = None, *, below_row=None,
            variant: str = "auto") -> torch.Tensor:
    """Filter a band of rows.  ``below_row`` is the row that follows the band: a
    tensor (w pixels), a raw device address (int, e.g. a peer-GPU pointer), or
    None when the band is the bottom of the image (clamp-to-edge)."""
    h, w = _as_image(img)
    if out is None:
        out = torch.empty_like(img)
    below = N.ptr(below_row) if below_row is not None else N.c_void_p(0)
    N.check(N.lib().b200_roberts(N.ptr(img), N.ptr(out), w, h, below, ROBERTS_VARIANTS[variant],
                                 N.current_stream_ptr()), "b200_roberts")
    return out


def roberts_geom(img: torch.Tensor, block: tuple[int, int], grid: tuple[int, int],
                 out: torch.Tensor | None = None) -> torch.Tensor:
    """Launch-geometry sweep entry (ref: lab2/src/to_plot.cu:57-64,106)."""
    h, w = _as_image(img)
    if out is None:
        out = torch.empty_like(img)
    N.check(N.lib().b200_roberts_geom(N.ptr(img), N.ptr(out), w, h, block[0], block[1], grid[0],
                                      grid[1], N.current_stream_ptr()), "b200_roberts_geom")
    return out
