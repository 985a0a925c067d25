"""lab3 pipeline (ref: lab3/src/main.cu:78-172).

The reference computes class means / covariances on the host and only classifies
on the GPU; here both stages run on the device: exact integer moments -> mean and
inverse covariance -> per-pixel argmin (direct, or through a 2^24-entry colour
table for large images).  ``stats='host'`` reproduces the reference's two-pass
host statistics in input order instead.
"""
from __future__ import annotations

import numpy as np
import torch

from .. import _native as N
from ..ops import classify as C
from ..utils.graphs import GraphedStep


class MahalanobisClassifier:
    def __init__(self, device: str | torch.device = "cuda", variant: str = "auto", stats: str = "device"):
        if stats not in ("device", "host"):
            raise ValueError("stats must be 'device' or 'host'")
        self.device, self.variant, self.stats = torch.device(device), variant, stats
        self._scratch = None

    def fit(self, img: torch.Tensor, classes) -> torch.Tensor:
        """Class parameters [nc, 12] from sample points [(np_c, 2) arrays of (x, y)]."""
        if self.stats == "host":
            host = img.cpu().numpy() if isinstance(img, torch.Tensor) else np.asarray(img)
            return torch.from_numpy(self.reference_params(host, classes, mode="ref")).to(self.device)
        return C.lab3_params_from_points(img, classes)

    def classify_(self, img: torch.Tensor, params: torch.Tensor) -> torch.Tensor:
        npix = img.shape[0] * img.shape[1]
        need = N.lib().b200_lab3_classify_scratch_bytes(npix, C.CLASSIFY_VARIANTS[self.variant])
        if need and (self._scratch is None or self._scratch.numel() < need):
            self._scratch = torch.empty(need, dtype=torch.uint8, device=self.device)
        return C.lab3_classify(img, params, variant=self.variant, scratch=self._scratch)

    def run_device(self, img: torch.Tensor, classes) -> torch.Tensor:
        return self.classify_(img, self.fit(img, classes))

    def graphed(self, img: torch.Tensor, classes) -> GraphedStep:
        """Statistics + classification of the fixed buffer ``img`` with fixed sample points as ONE
        CUDA-graph launch (moments, finalize, classify — plus table build and cell table for large
        images).  Copy a new image into ``img``, call the returned step; the sample points are
        uploaded here, outside the capture."""
        if self.stats != "device":
            raise ValueError("graph capture needs the device statistics path")
        pts, off = C.pack_points(classes, w=img.shape[1], h=img.shape[0])
        d_pts, d_off = torch.from_numpy(pts).to(img.device), torch.from_numpy(off).to(img.device)
        nc = len(classes)

        def step():
            return self.classify_(img, C.lab3_finalize(C.lab3_moments(img, d_pts, d_off, nc)))

        return GraphedStep(step, device=img.device)

    def run_host(self, host_img: torch.Tensor, classes, host_out: torch.Tensor | None = None):
        d = host_img.to(self.device, non_blocking=True)
        self.run_device(d, classes)
        if host_out is None:
            host_out = torch.empty_like(host_img, pin_memory=True)
        host_out.copy_(d, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return host_out

    # ----------------------------------------------------------------- oracle
    @staticmethod
    def reference_params(rgba: np.ndarray, classes, mode: str = "moments") -> np.ndarray:
        img = np.ascontiguousarray(rgba, dtype=np.uint8)
        h, w = img.shape[:2]
        pts, off = C.pack_points(classes, w=w, h=h)
        nc = len(classes)
        params = np.zeros((nc, C.N_PARAMS), dtype=np.float64)
        L = N.lib()
        if mode == "ref":
            L.b200_cpu_lab3_stats_ref(N.ptr(img), w, N.ptr(pts), N.ptr(off), nc, N.ptr(params))
        else:
            mom = np.zeros((nc, C.N_MOMENTS), dtype=np.int64)
            L.b200_cpu_lab3_moments(N.ptr(img), w, 0, h, N.ptr(pts), N.ptr(off), nc, N.ptr(mom))
            L.b200_cpu_lab3_finalize(N.ptr(mom), nc, N.ptr(params))
        return params

    @staticmethod
    def reference_classify(rgba: np.ndarray, params: np.ndarray) -> np.ndarray:
        out = np.array(rgba, dtype=np.uint8, copy=True, order="C")
        p = np.ascontiguousarray(params, dtype=np.float64)
        N.lib().b200_cpu_lab3_classify(N.ptr(out), out.shape[0] * out.shape[1], p.shape[0], N.ptr(p))
        return out
