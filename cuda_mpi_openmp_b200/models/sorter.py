"""lab5 pipeline: ascending sort of typed key arrays (file format of lab5/data/*)."""
from __future__ import annotations

import numpy as np
import torch

from .. import _native as N
from ..ops import sort as S
from ..utils.graphs import GraphedStep


class RadixSorter:
    def __init__(self, device: str | torch.device = "cuda"):
        self.device = torch.device(device)
        self._tmp = None
        self._scratch = None

    def _workspace(self, keys: torch.Tensor):
        if self._tmp is None or self._tmp.numel() < keys.numel() or self._tmp.dtype != keys.dtype:
            self._tmp = torch.empty_like(keys)
        need = S.sort_scratch_bytes(keys.numel(), keys.dtype)
        if self._scratch is None or self._scratch.numel() < need:
            self._scratch = torch.empty(need, dtype=torch.uint8, device=self.device)
        return self._tmp[: keys.numel()], self._scratch

    def run_device(self, keys: torch.Tensor) -> torch.Tensor:
        tmp, scratch = self._workspace(keys)
        return S.sort_(keys, tmp, scratch)

    def graphed(self, keys: torch.Tensor) -> GraphedStep:
        """The whole sort of the fixed buffer ``keys`` (memset, histogram, scan, four passes) as
        ONE CUDA-graph launch: copy new keys into ``keys``, call the returned step.  Pays for
        small arrays, where the seven launches cost more than the kernels."""
        tmp, scratch = self._workspace(keys)
        return GraphedStep(lambda: S.sort_(keys, tmp, scratch), device=keys.device)

    def run_host(self, host_keys: torch.Tensor, host_out: torch.Tensor | None = None):
        d = host_keys.to(self.device, non_blocking=True)
        self.run_device(d)
        if host_out is None:
            host_out = torch.empty_like(host_keys, pin_memory=True)
        host_out.copy_(d, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return host_out

    @staticmethod
    def reference(keys: np.ndarray) -> np.ndarray:
        kt = {np.dtype(np.int32): 0, np.dtype(np.float32): 1, np.dtype(np.uint8): 2,
              np.dtype(np.uint32): 3}[np.dtype(keys.dtype)]
        out = np.array(keys, copy=True, order="C")
        N.lib().b200_cpu_sort(N.ptr(out), out.size, kt)
        return out
