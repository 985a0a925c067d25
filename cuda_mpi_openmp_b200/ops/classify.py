"""lab3: Mahalanobis pixel classifier (ref: lab3/src/main.cu:40-76, :102-152)."""
from __future__ import annotations

import numpy as np
import torch

from .. import _native as N

CLASSIFY_VARIANTS = {"auto": 0, "direct": 1, "lut": 2}
MAX_CLASSES, N_MOMENTS, N_PARAMS = 32, 10, 12


def pack_points(classes, w: int | None = None, h: int | None = None) -> tuple[np.ndarray, np.ndarray]:
    """[(np_c, 2) int arrays of (x, y)] -> flat int32 pts, int32 offsets (nc + 1).

    With ``w``/``h`` (the WHOLE image, not a band) every point is checked against the image
    bounds here, once, on the host.  The kernels skip a point outside their band because
    another band owns it; a point outside the whole image would be skipped by every band
    and silently vanish from the statistics, so it is an error."""
    if not 1 <= len(classes) <= MAX_CLASSES:
        raise ValueError(f"class count must be 1..{MAX_CLASSES}")
    arrs = [np.asarray(c, dtype=np.int32).reshape(-1, 2) for c in classes]
    if w is not None and h is not None:
        for c, a in enumerate(arrs):
            bad = (a[:, 0] < 0) | (a[:, 0] >= w) | (a[:, 1] < 0) | (a[:, 1] >= h)
            if bad.any():
                x, y = a[np.argmax(bad)]
                raise ValueError(f"class {c}: sample point ({x}, {y}) lies outside the {w}x{h} image")
    off = np.zeros(len(arrs) + 1, dtype=np.int32)
    off[1:] = np.cumsum([a.shape[0] for a in arrs])
    pts = np.concatenate(arrs, axis=0).reshape(-1) if off[-1] else np.zeros(0, np.int32)
    return np.ascontiguousarray(pts, dtype=np.int32), off


def lab3_moments(img: torch.Tensor, pts: torch.Tensor, class_off: torch.Tensor, nc: int,
                 moments: torch.Tensor | None = None, *, y0: int = 0) -> torch.Tensor:
    """Exact int64 moments [nc, 10] of the sample pixels that fall in this band
    (rows y0 .. y0 + h).  Accumulates into ``moments`` when given."""
    h, w = img.shape[0], img.shape[1]
    if moments is None:
        moments = torch.zeros((nc, N_MOMENTS), dtype=torch.int64, device=img.device)
    N.check(N.lib().b200_lab3_moments(N.ptr(img), w, y0, h, N.ptr(pts), N.ptr(class_off), nc,
                                      N.ptr(moments), N.current_stream_ptr()), "b200_lab3_moments")
    return moments


def lab3_finalize(moments: torch.Tensor, params: torch.Tensor | None = None) -> torch.Tensor:
    """moments [nc, 10] -> params [nc, 12] (mean, inverse covariance) in float64."""
    nc = moments.shape[0]
    if params is None:
        params = torch.empty((nc, N_PARAMS), dtype=torch.float64, device=moments.device)
    N.check(N.lib().b200_lab3_finalize(N.ptr(moments), nc, N.ptr(params), N.current_stream_ptr()),
            "b200_lab3_finalize")
    return params


def lab3_params_from_points(img: torch.Tensor, classes) -> torch.Tensor:
    pts, off = pack_points(classes, w=img.shape[1], h=img.shape[0])
    d_pts = torch.from_numpy(pts).to(img.device)
    d_off = torch.from_numpy(off).to(img.device)
    return lab3_finalize(lab3_moments(img, d_pts, d_off, len(classes)))


def lab3_classify(img: torch.Tensor, params: torch.Tensor, *, variant: str = "auto",
                  scratch: torch.Tensor | None = None) -> torch.Tensor:
    """In place: alpha byte := index of the nearest class.  Returns ``img``."""
    npix = img.shape[0] * img.shape[1]
    nc = params.shape[0]
    v = CLASSIFY_VARIANTS[variant]
    need = N.lib().b200_lab3_classify_scratch_bytes(npix, v)
    if need and (scratch is None or scratch.numel() * scratch.element_size() < need):
        scratch = torch.empty(need, dtype=torch.uint8, device=img.device)
    sbytes = scratch.numel() * scratch.element_size() if scratch is not None else 0
    N.check(N.lib().b200_lab3_classify(N.ptr(img), npix, nc, N.ptr(params), v, N.ptr(scratch), sbytes,
                                       N.current_stream_ptr()), "b200_lab3_classify")
    return img


def lab3_classify_geom(img: torch.Tensor, params: torch.Tensor, blocks: int, threads: int):
    """<<<blocks, threads>>> sweep entry (ref: lab3/src/to_plot.cu:76-81,162)."""
    npix = img.shape[0] * img.shape[1]
    N.check(N.lib().b200_lab3_classify_geom(N.ptr(img), npix, params.shape[0], N.ptr(params), blocks,
                                            threads, N.current_stream_ptr()), "b200_lab3_classify_geom")
    return img
