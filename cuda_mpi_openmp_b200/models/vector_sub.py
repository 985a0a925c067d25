"""lab1 pipeline (ref: lab1/src/main.cu:36-89: read -> H2D -> kernel -> D2H -> print)."""
from __future__ import annotations

import numpy as np
import torch

from ..ops import vecsub


class VectorSub:
    def __init__(self, device: str | torch.device = "cuda"):
        self.device = torch.device(device)
        self._bufs: dict = {}

    def _buf(self, name: str, n: int) -> torch.Tensor:
        t = self._bufs.get(name)
        if t is None or t.numel() < n:
            t = torch.empty(n, dtype=torch.float64, device=self.device)
            self._bufs[name] = t
        return t[:n]

    def run_device(self, a: torch.Tensor, b: torch.Tensor, out: torch.Tensor | None = None):
        return vecsub(a, b, out)

    def run_host(self, a, b, out=None):
        """a, b: host float64 arrays/tensors (pinned for full PCIe speed)."""
        ta = torch.as_tensor(a, dtype=torch.float64)
        tb = torch.as_tensor(b, dtype=torch.float64)
        n = ta.numel()
        if out is None:
            out = torch.empty(n, dtype=torch.float64, pin_memory=torch.cuda.is_available())
        da, db, dc = self._buf("a", n), self._buf("b", n), self._buf("c", n)
        da.copy_(ta, non_blocking=True)
        db.copy_(tb, non_blocking=True)
        vecsub(da, db, dc)
        out.copy_(dc, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return out

    @staticmethod
    def reference(a, b):
        return np.asarray(a, dtype=np.float64) - np.asarray(b, dtype=np.float64)
