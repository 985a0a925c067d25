"""CUDA graphs for launch-bound sequences.

A sort of a few ten-thousand keys is seven launches (memset, histogram, scan, four
passes), a lab3 step on a harness-sized image is three; at those sizes the kernels take
microseconds and the time goes into launching them.  ``GraphedStep`` captures such a
sequence once — every libb200labs entry point launches on torch's current stream, so
the capture stream sees all of it — and replays it as one graph launch.

Only single-GPU sequences are captured: the fused multi-GPU kernels take the hand-shake
epoch as a kernel argument, which a replay would repeat.

Checked on hardware by tests/test_gpu_graphs.py (sort, lab3 step and a plain kernel: every replay
on new buffer contents equals the eager path, and goes through no C entry point).

(No counterpart in the reference: it launches one kernel per process.)
"""
from __future__ import annotations

from typing import Callable, Optional

import torch


class GraphedStep:
    """``step = GraphedStep(fn)`` runs ``fn`` a few times eagerly on a side stream (first calls
    set kernel attributes and size workspaces), captures one more call into a CUDA graph, and
    ``step()`` replays it on the current stream.  ``fn`` must use fixed buffers: a replay reads
    and writes the addresses seen during capture, so new inputs are copied INTO those buffers
    before ``step()``.  Tensors ``fn`` allocates during capture live in the graph's private pool
    and stay valid between replays; whatever ``fn`` returned at capture time is ``.result``."""

    def __init__(self, fn: Callable[[], object], warmup: int = 2, device: Optional[torch.device] = None):
        # the graph records raw addresses: everything fn's closure owns (workspaces, uploaded
        # tables, the model object) must live as long as the graph does
        self._fn = fn
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        side = torch.cuda.Stream(self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):
                fn()
        torch.cuda.current_stream(self.device).wait_stream(side)
        torch.cuda.synchronize(self.device)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph, stream=side):
            self.result = fn()

    def __call__(self):
        self.graph.replay()
        return self.result
