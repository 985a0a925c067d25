"""Experiment runner and verifier.

Contract kept from the reference (ref: tester.py):
* binaries are driven over stdin; if both kernel sizes are truthy, ``"k1\\nk2\\n"`` is
  prepended, a 2-list size becoming ``"a\\nb"`` (tester.py:113-121);
* line 0 of stdout carries ``execution time: <X ms>`` (tester.py:16,78-91), the
  rest goes to the processor's ``get_task_result`` / ``verify_result``;
* all tasks verified -> ``stats_<bin>.csv`` beside the binary, otherwise
  ``failed_<bin>.csv`` with the failing rows (tester.py:260-285);
* CUDA and optional CPU experiment, then ``median_execution_time.png``.

Differences by design: subprocesses are awaited through asyncio (bounded by
``max_parallel``, default 1 so timings do not overlap on the GPU), matplotlib is
optional (a Pillow chart, then a text table, are the fallbacks), and processors
receive ``device_info`` as a keyword they are free to ignore.
"""
from __future__ import annotations

import asyncio
import json
import os
import re
import time
import traceback
from dataclasses import asdict, dataclass
from typing import Any, Dict, List, Optional, Tuple

import numpy as np
import pandas as pd

TIME_KERNEL_EXE_PATTERN = r"execution time: <([\d.]+) ms>"
_TIME_RE = re.compile(TIME_KERNEL_EXE_PATTERN)


def parse_kernel_time(text: str) -> Optional[float]:
    m = _TIME_RE.search(text)
    return float(m.group(1)) if m else None


@dataclass
class TaskResult:
    test_verification_result: Optional[bool]
    task_result: Optional[Any]
    time_kernel_exe_ms: Optional[float]


@dataclass
class SubProcessResult(TaskResult):
    status: bool
    err: Optional[str]
    debug_data: Optional[Dict[str, Any]]


class BaseLabProcessor:
    """Plug-in that generates a binary's stdin and checks what comes back."""

    def __init__(self, seed: int = 42):
        self.seed = seed
        np.random.seed(seed)

    def get_attr(self) -> Dict[str, Any]:
        return {}

    async def pre_process(self, **kwargs) -> Tuple[str, Dict[str, Any], Dict[str, Any]]:
        raise NotImplementedError

    async def verify_result(self, task_result: Any, **kwargs) -> bool:
        raise NotImplementedError

    async def get_task_result(self, task_result_string: str, **kwargs) -> Any:
        raise NotImplementedError

    async def post_process(self, result_stdout: str, **kwargs) -> TaskResult:
        first, _, rest = result_stdout.partition("\n")
        task_result = await self.get_task_result(rest, **kwargs)
        ok = await self.verify_result(task_result, **kwargs)
        return TaskResult(test_verification_result=ok, task_result=task_result,
                          time_kernel_exe_ms=parse_kernel_time(first))


def _device_tag(binary_path: str, k1, k2) -> str:
    tag = f"{os.path.basename(binary_path)}_{k1}_{k2}"
    for old in (", ", " ", "[", "]"):
        tag = tag.replace(old, "_")
    return tag


def _size_to_stdin(k) -> str:
    return f"{k[0]}\n{k[1]}" if isinstance(k, (list, tuple)) else f"{k}"


async def run_subprocess(binary_path: str, lab_processor: BaseLabProcessor, return_inp: bool,
                         kernel_size_1=None, kernel_size_2=None,
                         gate: Optional[asyncio.Semaphore] = None) -> SubProcessResult:
    debug_data: Optional[Dict[str, Any]] = None

    def failed(err: str) -> SubProcessResult:
        return SubProcessResult(test_verification_result=None, task_result=None,
                                time_kernel_exe_ms=None, debug_data=debug_data, status=False, err=err)

    try:
        stdin_text, verify_ctx, debug_data = await lab_processor.pre_process(
            device_info=_device_tag(binary_path, kernel_size_1, kernel_size_2))
        if kernel_size_1 and kernel_size_2:
            stdin_text = f"{_size_to_stdin(kernel_size_1)}\n{_size_to_stdin(kernel_size_2)}\n{stdin_text}"
        if return_inp:
            debug_data["input_str"] = stdin_text
        async with (gate or asyncio.Semaphore(1)):
            proc = await asyncio.create_subprocess_exec(
                binary_path, stdin=asyncio.subprocess.PIPE, stdout=asyncio.subprocess.PIPE,
                stderr=asyncio.subprocess.PIPE)
            out, err = await proc.communicate(stdin_text.encode())
        if proc.returncode != 0:
            msg = err.decode(errors="replace").strip()
            print(f"[CUDA ERROR] {msg}")
            return failed(msg or f"exit code {proc.returncode}")
        res = await lab_processor.post_process(result_stdout=out.decode(errors="replace"), **verify_ctx)
        return SubProcessResult(test_verification_result=res.test_verification_result,
                                task_result=res.task_result, time_kernel_exe_ms=res.time_kernel_exe_ms,
                                debug_data=debug_data, status=True, err=None)
    except Exception:  # noqa: BLE001 - any processor/IO failure becomes a failed row
        traceback.print_exc()
        return failed(traceback.format_exc())


async def print_stat_time(values) -> None:
    v = np.asarray([x for x in values if x is not None], dtype=float)
    if v.size == 0:
        print("no timings")
        return
    print(f"Mean: {v.mean()} ms")
    print(f"Median: {np.median(v)} ms")
    print(f"Min: {v.min()} ms")
    print(f"Max: {v.max()} ms")
    print(f"Standard Deviation: {v.std()} ms")


class BaseTester:
    def __init__(self, binary_path_cuda: str, k_times: int, kernel_sizes: List[List],
                 metadata_columns2plot: List[str], binary_path_cpu: Optional[str] = None,
                 return_inp: bool = False, return_task_res: bool = False, max_parallel: int = 1):
        self.binary_path_cuda = binary_path_cuda
        self.binary_path_cpu = binary_path_cpu
        self.metadata_columns2plot = metadata_columns2plot or []
        self.dir2save = os.path.dirname(binary_path_cuda)
        self.kernel_sizes = kernel_sizes
        self.k_times = k_times
        self.return_inp = return_inp
        self.return_task_res = return_task_res
        self.max_parallel = max(1, int(max_parallel))

    async def run_experiment(self, binary_path: str, kernel_sizes: List[List],
                             lab_processor: BaseLabProcessor) -> pd.DataFrame:
        bin_name = os.path.splitext(os.path.basename(binary_path))[0]
        print(f"[Experiment bin_name=<{bin_name}>] START")
        gate = asyncio.Semaphore(self.max_parallel)
        rows: List[Dict[str, Any]] = []
        for i in range(self.k_times):
            for k1, k2 in kernel_sizes:
                print(f"[Experiment bin_name=<{bin_name}> task={i} kernel_size=<{[k1, k2]}>] started")
                t_created = time.time()
                # sequential await keeps inputs in round-robin order, like the reference
                result = await run_subprocess(binary_path, lab_processor, self.return_inp, k1, k2, gate)
                row = {"idx_run_time": i, "bin_name": bin_name, "kernel_size": [k1, k2]}
                row.update({k: v for k, v in asdict(result).items() if k != "debug_data"})
                row.update(lab_processor.get_attr() or {})
                if isinstance(result.debug_data, dict):
                    row.update(result.debug_data)
                row["time_exe_ms_from_start_run_time_bin_name"] = (time.time() - t_created) * 1000
                rows.append(row)
                print(f"[Experiment bin_name=<{bin_name}> task={i} kernel_size=<{[k1, k2]}>] "
                      f"finished with `time_kernel_exe_ms`: {row['time_kernel_exe_ms']} ms")

        drop = set() if self.return_task_res else {"task_result"}
        strip = lambda r: {k: v for k, v in r.items() if k not in drop}  # noqa: E731
        if rows and all(r.get("test_verification_result") for r in rows):
            await print_stat_time([r["time_kernel_exe_ms"] for r in rows])
            df = pd.DataFrame([strip(r) for r in rows])
            df.to_csv(os.path.join(self.dir2save, f"stats_{bin_name}.csv"), index=False)
            print(f"[Experiment bin_name=<{bin_name}>] SUCCESS!")
            return df
        bad = pd.DataFrame([strip(r) for r in rows if not r.get("test_verification_result")])
        print(f"[Experiment bin_name=<{bin_name}>] FAILED: len={bad.shape[0]}!")
        bad.to_csv(os.path.join(self.dir2save, f"failed_{bin_name}.csv"), index=False)
        return pd.DataFrame()

    async def run_experiments(self, lab_processor: BaseLabProcessor) -> Optional[pd.DataFrame]:
        t0 = time.time()
        print("[Experiments] START")
        df = await self.run_experiment(self.binary_path_cuda, self.kernel_sizes, lab_processor)
        df["device"] = ["CUDA"] * df.shape[0]
        if self.binary_path_cpu:
            cpu = await self.run_experiment(self.binary_path_cpu, [[None, None]], lab_processor)
            cpu["device"] = ["CPU"] * cpu.shape[0]
            df = pd.concat([df, cpu], ignore_index=True)
        if df.shape[0] > 0:
            self.plot_df_score(df_scores=df)
        print(f"[Experiments] FINISH time exe: {time.time() - t0}")
        return df

    # ------------------------------------------------------------------ chart
    def plot_df_score(self, df_scores: pd.DataFrame) -> str:
        from .plotting import median_bar_chart

        df = df_scores.copy()
        df["kernel_size"] = df["kernel_size"].apply(json.dumps)
        grouped = (df.groupby(["device", "kernel_size"])
                   .agg(median_time=("time_kernel_exe_ms", "median"),
                        sample_count=("time_kernel_exe_ms", "size")).reset_index())
        grouped["label"] = [d if d == "CPU" else f"{d}_{k}"
                            for d, k in zip(grouped["device"], grouped["kernel_size"])]
        legend = ""
        for col in self.metadata_columns2plot:
            if col in df.columns:
                legend += f"{col}: [{', '.join(map(str, df[col].unique()))}]\n"
        legend += "\nSample Count by Group:\n"
        for _, r in grouped.iterrows():
            legend += f"{r['label']}: {r['sample_count']} samples\n"
        path = os.path.join(self.dir2save, "median_execution_time.png")
        return median_bar_chart(list(grouped["label"]), list(grouped["median_time"]), legend, path)
