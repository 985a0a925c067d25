"""Device-side timing helpers (CUDA events on the launching stream)."""
from __future__ import annotations

from typing import Callable, List

import torch

from .. import _native as N


class L2Flusher:
    """Writes a buffer larger than the 126 MB L2 so the next kernel starts cold."""

    def __init__(self, device=None, nbytes: int = 256 << 20):
        self.buf = torch.empty(nbytes, dtype=torch.uint8, device=device or "cuda")

    def __call__(self) -> None:
        N.check(N.lib().b200_flush_l2(N.ptr(self.buf), self.buf.numel(), N.current_stream_ptr()),
                "b200_flush_l2")


def cuda_time_ms(fn: Callable[[], None], iters: int = 10, warmup: int = 3,
                 flush: L2Flusher | None = None) -> List[float]:
    """Per-iteration device milliseconds of ``fn`` (one event pair per iteration,
    synchronised on both sides; optional L2 flush between iterations)."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    out = []
    for _ in range(iters):
        if flush is not None:
            flush()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        out.append(a.elapsed_time(b))
    return out
