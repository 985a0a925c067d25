#!/usr/bin/env python
"""lab2 TMA kernel on the 65536^2 image: does the buffer layout or the launch footprint matter?

  * input and output as two allocations, or carved from one buffer with a pad between them;
  * the whole image in one launch, or as K launches over K row bands (halo row = first row of
    the next band), timed as one region;
  * a plain device copy of the same bytes beside each, as the yardstick.
Prints one JSON line per configuration (device ms, median of --iters)."""
import argparse
import json
import statistics
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from cuda_mpi_openmp_b200 import ops  # noqa: E402
from cuda_mpi_openmp_b200.utils import cuda_time_ms  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--h", type=int, default=65536)
    ap.add_argument("--w", type=int, default=65536)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--pads", default="sep,0,4096,1052672,34603008")
    ap.add_argument("--bands", default="1,2,4,8")
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    h, w = args.h, args.w
    nbytes = h * w * 4

    def fill(img):
        g = torch.Generator(device=dev).manual_seed(3)
        flat = img.view(-1)
        for o in range(0, flat.numel(), 1 << 28):
            part = flat[o:o + (1 << 28)]
            part.copy_(torch.randint(-2 ** 31, 2 ** 31 - 1, (part.numel(),), dtype=torch.int32, device=dev,
                                     generator=g))

    def report(tag, img, out, bands):
        rows = h // bands

        def run():
            for k in range(bands):
                a, b = k * rows, (k + 1) * rows
                below = img[b] if b < h else None
                ops.roberts(img[a:b], out[a:b], below_row=below, variant="tma")

        ms = statistics.median(cuda_time_ms(run, iters=args.iters, warmup=2))
        cp = statistics.median(cuda_time_ms(lambda: out.copy_(img), iters=args.iters, warmup=1))
        print(json.dumps({"layout": tag, "bands": bands, "h": h, "w": w, "ms": round(ms, 4),
                          "GBps": round(2 * nbytes / ms / 1e6, 1), "copy_ms": round(cp, 4),
                          "copy_GBps": round(2 * nbytes / cp / 1e6, 1),
                          "out_minus_in_bytes": out.data_ptr() - img.data_ptr()}), flush=True)

    for pad in args.pads.split(","):
        if pad == "sep":
            img = torch.empty((h, w), dtype=torch.int32, device=dev)
            out = torch.empty_like(img)
            buf = None
        else:
            buf = torch.empty(2 * nbytes + int(pad), dtype=torch.uint8, device=dev)
            img = buf[:nbytes].view(torch.int32).view(h, w)
            out = buf[nbytes + int(pad):][:nbytes].view(torch.int32).view(h, w)
        fill(img)
        for bands in ([int(b) for b in args.bands.split(",")] if pad in ("sep", "0") else [1]):
            report(f"pad {pad}", img, out, bands)
        del img, out, buf
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
