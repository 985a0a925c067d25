"""Command-line front end of the harness (ref: run_test.py:18-83).

    python run_test.py --binary_path_cuda ./lab2/src/to_plot_cuda_exe \
        [--binary_path_cpu P] [--k_times N] [--kernel_sizes JSON] \
        [--metadata_columns2plot JSON] [--return_inp] [--return_task_res] [--<any> v ...]

The lab is the name of the binary's grandparent directory (``<labN>/src/<exe>``);
unknown ``--key value`` pairs go to the lab processor's constructor.
"""
from __future__ import annotations

import argparse
import asyncio
import json
import os
from typing import List, Optional

from .argkw import unknown_args_to_kwargs
from .core import BaseTester
from .processors import LAB_PROCESSORS


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description="Black-box timing and verification of a lab binary (one subprocess per run).")
    p.add_argument("--binary_path_cuda", type=str, required=True, help="CUDA build to test, at <lab>/<dir>/<exe> (the lab is read from the path).")
    p.add_argument("--binary_path_cpu", type=str, default=None, help="Optional CPU build, timed beside the CUDA one.")
    p.add_argument("--k_times", type=int, default=20, help="Runs per launch geometry.")
    p.add_argument("--return_inp", action="store_true", help="Keep each run's stdin text in the result rows.")
    p.add_argument("--return_task_res", action="store_true", help="Keep each run's parsed output in the result rows.")
    p.add_argument("--kernel_sizes", type=str, default="[[512, 512]]",
                   help="Launch geometries to sweep, JSON list of [size1, size2] (ints or 2-lists)")
    p.add_argument("--metadata_columns2plot", type=str, default="[]",
                   help='Columns listed in the chart legend, JSON, e.g. \'["filename"]\'')
    p.add_argument("--max_parallel", type=int, default=1,
                   help="Subprocesses allowed in flight (1 keeps GPU timings clean).")
    return p


def lab_name_of(binary_path: str) -> str:
    return os.path.basename(os.path.dirname(os.path.dirname(os.path.abspath(binary_path))))


def main(argv: Optional[List[str]] = None):
    args, unknown = build_parser().parse_known_args(argv)
    kwargs = unknown_args_to_kwargs(unknown)
    kernel_sizes = json.loads(args.kernel_sizes) if args.kernel_sizes else [[None, None]]
    columns = json.loads(args.metadata_columns2plot) if args.metadata_columns2plot else []
    lab = lab_name_of(args.binary_path_cuda)
    if lab not in LAB_PROCESSORS:
        raise SystemExit(f"cannot tell the lab from '{args.binary_path_cuda}': expected "
                         f"<lab>/<dir>/<exe> with <lab> in {sorted(LAB_PROCESSORS)}")
    print("Params:")
    print(f"return_inp=<{args.return_inp}>")
    print(f"return_task_res=<{args.return_task_res}>")
    print(f"lab_name=<{lab}>")
    print(f"binary_path_cuda=<{args.binary_path_cuda}>")
    print(f"binary_path_cpu=<{args.binary_path_cpu}>")
    print(f"k_times=<{args.k_times}>")
    print(f"kernel_sizes=<{kernel_sizes}>")
    print(f"kwargs=<{json.dumps(kwargs, indent=2)}>")
    print(f"metadata_columns2plot=<{json.dumps(columns, indent=2)}>")
    tester = BaseTester(binary_path_cuda=args.binary_path_cuda, binary_path_cpu=args.binary_path_cpu,
                        k_times=args.k_times, kernel_sizes=kernel_sizes, return_inp=args.return_inp,
                        return_task_res=args.return_task_res, metadata_columns2plot=columns,
                        max_parallel=args.max_parallel)
    processor = LAB_PROCESSORS[lab](**kwargs)
    return asyncio.run(tester.run_experiments(lab_processor=processor))


if __name__ == "__main__":
    main()
