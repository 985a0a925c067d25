"""Output checks that bench.py runs OUTSIDE its timed regions (its "verify" key), so that a
benchmark line is also evidence that the timed kernels produced the right bytes at the
benchmark's own size.  Every function works on one rank's shard; the caller reduces the
verdicts over the ranks."""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch


def xor_reduce(t: torch.Tensor) -> int:
    """XOR of all elements of an int32 tensor (torch has no xor reduction: fold halves)."""
    t = t.reshape(-1).view(torch.int32)
    if t.numel() == 0:
        return 0
    while t.numel() > 1:
        n = t.numel()
        half = n // 2
        folded = t[:half] ^ t[half:2 * half]
        if n & 1:
            folded[0] ^= t[n - 1]
        t = folded
    return int(t.item()) & 0xFFFFFFFF


def checksum(t: torch.Tensor) -> tuple[int, int]:
    """(sum as int64, xor) of a 32-bit tensor: equal multisets have equal checksums."""
    v = t.reshape(-1).view(torch.int32)
    return int(v.sum(dtype=torch.int64).item()), xor_reduce(v)


def is_sorted(t: torch.Tensor, chunk: int = 1 << 28) -> bool:
    """Non-decreasing, checked on the device in pieces (no full-size temporaries)."""
    n = t.numel()
    for o in range(0, max(n - 1, 0), chunk):
        e = min(n - 1, o + chunk)
        if not bool((t[o + 1:e + 1] >= t[o:e]).all().item()):
            return False
    return True


def roberts_blocks(band: torch.Tensor, out: torch.Tensor, below_row: Optional[torch.Tensor], y_starts,
                   rows: int = 64) -> dict:
    """Sampled row blocks of a filtered band against the CPU oracle.  band/out: int32 [h, w]
    device tensors; below_row: the row after the band (the neighbour GPU's first row) or None
    when the band ends the image.  A block's oracle input is its rows plus the row below."""
    from ..models import RobertsFilter

    h, w = band.shape
    bad, checked, used_halo = [], 0, False
    for y0 in y_starts:
        y0 = int(max(0, min(y0, h - 1)))
        y1 = min(h, y0 + rows)
        if y1 < h:
            src = band[y0:y1 + 1]
        elif below_row is not None:
            src = torch.cat([band[y0:y1], below_row.reshape(1, w).to(band.device)], dim=0)
            used_halo = True
        else:
            src = band[y0:y1]                      # true bottom of the image: the oracle clamps there too
        host = src.contiguous().cpu().numpy().view(np.uint8).reshape(src.shape[0], w, 4)
        want = RobertsFilter.reference(host)[: y1 - y0]
        got = out[y0:y1].cpu().numpy().view(np.uint8).reshape(y1 - y0, w, 4)
        checked += y1 - y0
        if not np.array_equal(want, got):
            bad.append(y0)
    return {"blocks": len(list(y_starts)), "rows_checked": checked, "bad_blocks": bad[:8],
            "halo_block_checked": used_halo, "ok": not bad}


def sample_block_starts(h: int, w: int, n_blocks: int = 16, rows: int = 64, seed: int = 0) -> list:
    """Block starts that always include row 0, the last rows of the band, the first block whose
    byte offset is >= 2^32 (if the band is that large) and otherwise spread pseudo-randomly."""
    ys = {0, max(0, h - rows)}
    first_4g = (1 << 32) // (4 * w) + 1
    if first_4g + rows <= h:
        ys.add(first_4g)
        ys.add(h - rows - 1 if h - rows - 1 > first_4g else first_4g)
    rng = np.random.default_rng(seed)
    while len(ys) < n_blocks and len(ys) < max(1, h // rows):
        ys.add(int(rng.integers(0, max(1, h - rows))))
    return sorted(ys)
