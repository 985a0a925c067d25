"""lab2 pipeline (ref: lab2/src/main.cu:54-126: read -> texture upload -> kernel -> D2H -> write).

``run_device`` filters an image that already lives on the GPU.  ``run_host`` is the
end-to-end path: the image sits in (pinned) host memory, is streamed through the
GPU in row chunks on several CUDA streams so that H2D copy, stencil kernel and D2H
copy of different chunks overlap, and lands in a host output buffer.  Each chunk
carries one extra row so the stencil's ``y+1`` look-ahead never leaves the chunk.
"""
from __future__ import annotations

import numpy as np
import torch

from .. import _native as N
from ..ops import roberts
from ..ops.roberts import ROBERTS_VARIANTS


class RobertsFilter:
    def __init__(self, device: str | torch.device = "cuda", variant: str = "auto",
                 chunk_bytes: int = 256 << 20, n_streams: int = 3):
        self.device = torch.device(device)
        self.variant = variant
        self.chunk_bytes = chunk_bytes
        self.n_streams = n_streams
        self._streams = None
        self._stage: dict = {}

    # ------------------------------------------------------------------ device
    def run_device(self, img: torch.Tensor, out: torch.Tensor | None = None, below_row=None):
        return roberts(img, out, below_row=below_row, variant=self.variant)

    # -------------------------------------------------------------------- host
    def _ensure_stage(self, rows: int, w: int):
        key = (rows, w)
        if self._stage.get("key") != key:
            self._stage = {
                "key": key,
                "in": [torch.empty((rows + 1, w), dtype=torch.int32, device=self.device)
                       for _ in range(self.n_streams)],
                "out": [torch.empty((rows, w), dtype=torch.int32, device=self.device)
                        for _ in range(self.n_streams)],
            }
        if self._streams is None:
            self._streams = [torch.cuda.Stream(device=self.device) for _ in range(self.n_streams)]

    def run_host(self, host_in: torch.Tensor, host_out: torch.Tensor | None = None,
                 below_row_host: torch.Tensor | None = None) -> torch.Tensor:
        """host_in: int32 [h, w] (or uint8 [h, w, 4]) CPU tensor, ideally pinned.
        ``below_row_host`` is the row after the band when this band is one shard of a
        taller image (None = clamp at the bottom)."""
        src = host_in.view(torch.int32).reshape(host_in.shape[0], host_in.shape[1])
        h, w = src.shape
        if host_out is None:
            host_out = torch.empty((h, w), dtype=torch.int32, pin_memory=True)
        dst = host_out.view(torch.int32).reshape(h, w)
        rows = max(1, min(h, self.chunk_bytes // (4 * w)))
        self._ensure_stage(rows, w)
        lib, variant = N.lib(), ROBERTS_VARIANTS[self.variant]
        cur = torch.cuda.current_stream(self.device)
        for s in self._streams:
            s.wait_stream(cur)
        for k, r0 in enumerate(range(0, h, rows)):
            r1 = min(h, r0 + rows)
            s = self._streams[k % self.n_streams]
            d_in, d_out = self._stage["in"][k % self.n_streams], self._stage["out"][k % self.n_streams]
            with torch.cuda.stream(s):
                n = r1 - r0
                if r1 < h:  # halo row comes from the next chunk of the same image
                    d_in[: n + 1].copy_(src[r0: r1 + 1], non_blocking=True)
                    below = d_in[n]
                else:
                    d_in[:n].copy_(src[r0:r1], non_blocking=True)
                    below = None
                    if below_row_host is not None:
                        d_in[n].copy_(below_row_host.view(torch.int32).reshape(w), non_blocking=True)
                        below = d_in[n]
                N.check(lib.b200_roberts(N.ptr(d_in), N.ptr(d_out), w, n,
                                         N.ptr(below) if below is not None else N.c_void_p(0),
                                         variant, N.c_void_p(s.cuda_stream)), "b200_roberts")
                dst[r0:r1].copy_(d_out[:n], non_blocking=True)
        for s in self._streams:
            cur.wait_stream(s)
        cur.synchronize()
        return host_out

    # ----------------------------------------------------------------- oracle
    @staticmethod
    def reference(rgba: np.ndarray) -> np.ndarray:
        """CPU oracle (FMA-faithful C++ restatement), uint8 [h, w, 4] -> same."""
        src = np.ascontiguousarray(rgba, dtype=np.uint8)
        out = np.empty_like(src)
        N.lib().b200_cpu_roberts(N.ptr(src), N.ptr(out), src.shape[1], src.shape[0])
        return out
