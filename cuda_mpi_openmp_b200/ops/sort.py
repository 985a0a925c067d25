"""lab5: in-place ascending radix sort (no reference code exists; see csrc/lab5_sort.cu)."""
from __future__ import annotations

import torch

from .. import _native as N

KEY_TYPES = {torch.int32: 0, torch.float32: 1, torch.uint8: 2, torch.uint32: 3}


def sort_scratch_bytes(n: int, dtype: torch.dtype) -> int:
    return int(N.lib().b200_sort_scratch_bytes(n, KEY_TYPES[dtype]))


def sort_(keys: torch.Tensor, tmp: torch.Tensor | None = None,
          scratch: torch.Tensor | None = None) -> torch.Tensor:
    """Sort a 1-D CUDA tensor in place (int32 / uint32 / float32 / uint8).
    ``tmp`` (same shape) and ``scratch`` (uint8, sort_scratch_bytes) may be
    passed in to keep allocations out of a timed region."""
    if keys.dim() != 1 or not keys.is_cuda or not keys.is_contiguous():
        raise ValueError("sort_ needs a contiguous 1-D CUDA tensor")
    kt = KEY_TYPES.get(keys.dtype)
    if kt is None:
        raise TypeError(f"unsupported key dtype {keys.dtype}")
    n = keys.numel()
    if tmp is None:
        tmp = torch.empty_like(keys)
    need = sort_scratch_bytes(n, keys.dtype)
    if scratch is None or scratch.numel() < need:
        scratch = torch.empty(need, dtype=torch.uint8, device=keys.device)
    N.check(N.lib().b200_sort(N.ptr(keys), N.ptr(tmp), n, kt, N.ptr(scratch), scratch.numel(),
                              N.current_stream_ptr()), "b200_sort")
    return keys
