"""Per-lab input generators / verifiers for the harness.

Parity with the reference processors (stdin layout, defaults, verification rule):
* lab1 — ref: lab1/lab1_processor.py:27-70.  The reference disables its check and,
  with the current tester, crashes on the ``device_info`` keyword; here the keyword
  is accepted and the result IS checked, against ``a - b`` recomputed from the
  *printed* inputs at relative 1e-10 (the printed text is what the binary sees).
* lab2 — ref: lab2/lab2_processor.py:16-172: fixed file list, round-robin inputs,
  exact hex equality against ``<data>_out_gt/<name>.*`` when a golden exists.
* lab3 — ref: lab3/lab3_processor.py:54-233: as lab2, plus the class definitions on
  stdin; every image uses the two hard-coded 4-point classes unless
  ``count_classes``/``count_pts`` ask for random ones.
* lab5 — new (the reference has no lab5 code): sorts ``int32 n + keys`` files,
  verified against ``numpy.sort``.
"""
from __future__ import annotations

import asyncio
import os
import shutil
import struct
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Tuple
from uuid import uuid4

import numpy as np

from ..utils import ImgData, download_file
from .core import BaseLabProcessor

NEW_LINE = "\n"
MAX_CLASSES = 32
MAX_NUM_POINTS = 2 ** 19

LAB2_FILENAMES = [
    "stalker2.png", "98.data", "AoE.png", "doom.png", "hf2.png", "starcraft.png", "warcraft.png",
    "test_01.txt", "test_02.txt", "lenna.png", "57.data", "95.data", "99.data", "02.data",
    "96.data", "97.data",
]
LAB3_FILENAMES = LAB2_FILENAMES + ["04.data", "09.data", "test_01_lab3.txt", "test_02_lab3.txt"]


@dataclass
class GroundTruthClass:
    lbl: int
    definition_points: np.ndarray  # (np, 2) integer (x, y) pairs


# the worked example of the lab3 task statement (ref: lab3/lab3_processor.py:42-51)
DEFAULT_CLASSES = [
    GroundTruthClass(0, np.array([[1, 2], [1, 0], [2, 2], [2, 1]])),
    GroundTruthClass(1, np.array([[0, 0], [0, 1], [1, 1], [2, 0]])),
]


def get_random_pts(w: int, h: int, count_pts: Optional[int] = None) -> np.ndarray:
    """Random sample points inside a random top-left sub-rectangle
    (ref: lab3/img_data_classifier.py:16-24)."""
    if count_pts is None:
        count_pts = int(np.random.randint(2, MAX_CLASSES + 2))
    x_hi = int(np.random.randint(1, w + 1))
    y_hi = int(np.random.randint(1, h + 1))
    return np.stack([np.random.randint(x_hi, size=count_pts),
                     np.random.randint(y_hi, size=count_pts)], axis=1)


def _fmt_vector(v: np.ndarray, precision: int) -> str:
    return np.array2string(v, separator=" ", max_line_width=np.inf, precision=precision,
                           threshold=np.iinfo(np.int64).max)[1:-1].strip()


# --------------------------------------------------------------------------- lab1
class Lab1Processor(BaseLabProcessor):
    def __init__(self, seed: int = 42, min_vector_size: int = 1024, max_vector_size: int = 3072,
                 atol: float = 1e-10, precision_array: int = 10, verify: bool = True, **_ignored):
        super().__init__(seed=seed)
        self.double_left, self.double_right = -1e100, 1e100
        self.min_vector_size, self.max_vector_size = min_vector_size, max_vector_size
        self.precision_array, self.atol, self.verify = precision_array, atol, verify

    async def pre_process(self, **kwargs):
        n = int(np.random.randint(self.min_vector_size, self.max_vector_size))
        a = np.random.uniform(self.double_left, self.double_right, n)
        b = np.random.uniform(self.double_left, self.double_right, n)
        a_txt, b_txt = _fmt_vector(a, self.precision_array), _fmt_vector(b, self.precision_array)
        ctx = {"first_vector": a, "second_vector": b,
               # what the binary actually parses: the 10-digit text, not the doubles
               "first_vector_printed": np.array(a_txt.split(), dtype=np.float64),
               "second_vector_printed": np.array(b_txt.split(), dtype=np.float64)}
        return f"{n}\n{a_txt}\n{b_txt}", ctx, {"vector_size": n}

    async def verify_result(self, task_result: np.ndarray, **kwargs) -> bool:
        if not self.verify:
            return True
        want = kwargs["first_vector_printed"] - kwargs["second_vector_printed"]
        if task_result.shape != want.shape:
            return False
        # the output is printed with %.10e: relative agreement, not absolute
        return bool(np.allclose(task_result, want, rtol=max(self.atol, 1e-10), atol=0.0))

    async def get_task_result(self, task_result_string: str, **kwargs) -> np.ndarray:
        return np.array(task_result_string.split(), dtype=np.float64)

    def get_attr(self):
        return {"min_vector_size": self.min_vector_size, "max_vector_size": self.max_vector_size,
                "atol": self.atol}


# ------------------------------------------------------------------ image-file labs
class _ImageLabProcessor(BaseLabProcessor):
    """Shared machinery of lab2/lab3: file discovery, golden lookup, round-robin."""

    FILENAMES: List[str] = []
    DEFAULT_DIR = "."

    def __init__(self, seed: int = 42, atol: float = 1e-10, precision_array: int = 10,
                 extra_links_to_png: Optional[List[str]] = None, dir_to_data: Optional[str] = None,
                 dir_to_data_out: Optional[str] = None, dir_to_data_out_gt: Optional[str] = None,
                 **_ignored):
        super().__init__(seed=seed)
        self.precision_array, self.atol = precision_array, atol
        dir_to_data = dir_to_data or self.DEFAULT_DIR
        parent, base = os.path.dirname(dir_to_data), os.path.basename(dir_to_data)
        dir_to_data_out_gt = dir_to_data_out_gt or os.path.join(parent, f"{base}_out_gt")
        self.dir_to_data_out = dir_to_data_out or os.path.join(parent, f"{base}_out")

        paths = [os.path.join(dir_to_data, fn) for fn in self.FILENAMES
                 if os.path.exists(os.path.join(dir_to_data, fn))]
        for link in (extra_links_to_png or []):
            paths.append(download_file(link, save_dir=dir_to_data, filename=f"{uuid4()}.png"))
        if not paths:
            raise FileNotFoundError(f"no known input images in {dir_to_data}")

        # start from an empty output tree; converted inputs (.data/.txt/.png siblings)
        # go under it instead of into the input directory, which stays untouched
        shutil.rmtree(self.dir_to_data_out, ignore_errors=True)
        cache = os.path.join(self.dir_to_data_out, "_inputs")
        os.makedirs(cache, exist_ok=True)

        self.data_input: Dict[int, Any] = {}
        self.data_output_gt: Dict[int, ImgData] = {}
        for i, path in enumerate(paths):
            item = ImgData(path2data=path, idx=i, out_dir=cache)
            self.data_input[i] = self._wrap_input(item)
            for ext in ("txt", "data", "png"):
                gt = os.path.join(dir_to_data_out_gt, f"{item.data_name}.{ext}")
                if os.path.exists(gt):
                    self.data_output_gt[i] = ImgData(path2data=gt, idx=i,
                                                     out_dir=os.path.join(cache, "gt"))
                    break
        self.current_index = 0
        self.lock = asyncio.Lock()

    def _wrap_input(self, item: ImgData):
        return item

    def _image_of(self, entry) -> ImgData:
        return entry

    async def get_next_item(self):
        async with self.lock:
            entry = self.data_input[self.current_index]
            self.current_index = (self.current_index + 1) % len(self.data_input)
            return entry

    def _out_path(self, device_info: Optional[str], item: ImgData) -> str:
        d = os.path.join(self.dir_to_data_out, device_info or "run")
        os.makedirs(d, exist_ok=True)
        return os.path.join(d, f"{item.data_name}.data")

    async def verify_result(self, task_result: ImgData, **kwargs) -> bool:
        idx = kwargs.get("idx_data")
        task_result.idx = idx
        gt = self.data_output_gt.get(idx)
        if gt is None:
            return True  # no golden for this input: unverified pass, like the reference
        norm = lambda s: "".join(s.split()).upper()  # noqa: E731
        ok = norm(task_result.hex) == norm(gt.hex)
        if not ok:
            src = self._image_of(self.data_input[idx])
            print(f"[verify_result] FAILED `verify_result`: `{task_result.data_name}`!")
            print(f"[verify_result] [input_data.hex] {src.hex.replace(NEW_LINE, ' ').upper()}")
            print(f"[verify_result] [task_result.hex] {task_result.hex.replace(NEW_LINE, ' ').upper()}")
            print(f"[verify_result] [ground_truth.hex] {gt.hex.replace(NEW_LINE, ' ').upper()}")
            print("[verify_result] SHOULD BE `task_result.hex` == `ground_truth.hex`!!!")
        return ok

    async def get_task_result(self, task_result_string: str, **kwargs) -> ImgData:
        return ImgData(path2data=kwargs["out_path_res"])

    def get_attr(self):
        return {"precision_array": self.precision_array, "atol": self.atol}


class Lab2Processor(_ImageLabProcessor):
    FILENAMES = LAB2_FILENAMES
    DEFAULT_DIR = "./lab2/data"

    async def pre_process(self, **kwargs):
        item: ImgData = await self.get_next_item()
        out_path = self._out_path(kwargs.get("device_info"), item)
        return (f"{item.c_data_bytes_path}\n{out_path}",
                {"idx_data": item.idx, "out_path_res": out_path},
                {"filename": f"{item.data_name}{item.data_ext} ({item.size:.5f} KB)"})


class Lab3Processor(_ImageLabProcessor):
    FILENAMES = LAB3_FILENAMES
    DEFAULT_DIR = "./lab3/data"

    def __init__(self, count_classes: Optional[int] = None, count_pts: Optional[int] = None, **kw):
        self.count_classes, self.count_pts = count_classes, count_pts
        super().__init__(**kw)

    def _wrap_input(self, item: ImgData) -> Tuple[ImgData, List[GroundTruthClass]]:
        if self.count_classes is None:
            classes = DEFAULT_CLASSES
        else:
            h, w = item.shape
            classes = [GroundTruthClass(i, get_random_pts(w, h, self.count_pts))
                       for i in range(int(self.count_classes))]
        if not 0 < len(classes) <= MAX_CLASSES:
            raise ValueError(f"class count {len(classes)} outside 1..{MAX_CLASSES}")
        h, w = item.shape
        for c in classes:
            p = c.definition_points
            if p.shape[0] > MAX_NUM_POINTS:
                raise ValueError("too many definition points")
            if (p[:, 0] >= w).any() or (p[:, 1] >= h).any():
                raise ValueError(f"{item.data_name}: class {c.lbl} has points outside the {w}x{h} image")
        return item, classes

    def _image_of(self, entry) -> ImgData:
        return entry[0]

    async def pre_process(self, **kwargs):
        item, classes = await self.get_next_item()
        out_path = self._out_path(kwargs.get("device_info"), item)
        rows = [f"{c.definition_points.shape[0]} {_fmt_vector(c.definition_points.reshape(-1), self.precision_array)}"
                for c in classes]
        stdin = f"{item.c_data_bytes_path}\n{out_path}\n{len(classes)}\n" + "\n".join(rows)
        sizes = {c.definition_points.shape[0] for c in classes}
        return (stdin, {"idx_data": item.idx, "out_path_res": out_path},
                {"filename": f"{item.data_name}{item.data_ext} ({item.size:.5f} KB)",
                 "stat_init_pts": f"Count Classes: {len(classes)}, Count Points: {sizes}"})


# --------------------------------------------------------------------------- lab5
_LAB5_DTYPES = {"int": np.int32, "uint": np.uint32, "float": np.float32, "uchar": np.uint8}


def _lab5_type_of(name: str) -> str:
    for key in ("float", "uchar", "uint", "int"):
        if key in name:
            return key
    return "int"


def read_array_file(path: str, type_name: str) -> np.ndarray:
    with open(path, "rb") as f:
        blob = f.read()
    (n,) = struct.unpack_from("<i", blob, 0)
    return np.frombuffer(blob, dtype=_LAB5_DTYPES[type_name], count=n, offset=4)


def write_array_file(path: str, arr: np.ndarray) -> None:
    with open(path, "wb") as f:
        f.write(struct.pack("<i", arr.size))
        f.write(np.ascontiguousarray(arr).tobytes())


class Lab5Processor(BaseLabProcessor):
    """Feeds ``lab5/src/to_plot_*`` with array files: the fixtures in ``dir_to_data``
    plus ``extra_random`` generated arrays of ``random_size`` keys per type."""

    def __init__(self, seed: int = 42, dir_to_data: str = "./lab5/data", dir_to_data_out: Optional[str] = None,
                 extra_random: int = 3, random_size: int = 1 << 16, **_ignored):
        super().__init__(seed=seed)
        parent, base = os.path.dirname(dir_to_data), os.path.basename(dir_to_data)
        self.dir_to_data_out = dir_to_data_out or os.path.join(parent, f"{base}_out")
        shutil.rmtree(self.dir_to_data_out, ignore_errors=True)
        os.makedirs(self.dir_to_data_out, exist_ok=True)
        self.items: List[Tuple[str, str]] = []  # (path, type)
        for fn in sorted(os.listdir(dir_to_data)):
            p = os.path.join(dir_to_data, fn)
            if os.path.isfile(p) and "." not in fn:
                self.items.append((p, _lab5_type_of(fn)))
        gen_dir = os.path.join(self.dir_to_data_out, "_generated")
        os.makedirs(gen_dir, exist_ok=True)
        rng = np.random.default_rng(seed)
        for i in range(int(extra_random)):
            for t in ("int", "float", "uchar"):
                if t == "int":
                    arr = rng.integers(-2 ** 31, 2 ** 31 - 1, size=random_size, dtype=np.int64).astype(np.int32)
                elif t == "float":
                    arr = (rng.standard_normal(random_size) * 1e3).astype(np.float32)
                else:
                    arr = rng.integers(0, 256, size=random_size, dtype=np.int64).astype(np.uint8)
                p = os.path.join(gen_dir, f"{t}_rand{i}_{random_size}")
                write_array_file(p, arr)
                self.items.append((p, t))
        if not self.items:
            raise FileNotFoundError(f"no array files in {dir_to_data}")
        self.current_index = 0
        self.lock = asyncio.Lock()

    async def pre_process(self, **kwargs):
        async with self.lock:
            idx = self.current_index
            self.current_index = (idx + 1) % len(self.items)
        path, t = self.items[idx]
        d = os.path.join(self.dir_to_data_out, kwargs.get("device_info") or "run")
        os.makedirs(d, exist_ok=True)
        out_path = os.path.join(d, os.path.basename(path) + ".sorted")
        n = read_array_file(path, t).size
        return (f"{path}\n{out_path}\n{t}", {"idx_data": idx, "out_path_res": out_path},
                {"filename": f"{os.path.basename(path)} ({n} keys, {t})"})

    async def get_task_result(self, task_result_string: str, **kwargs) -> np.ndarray:
        _, t = self.items[kwargs["idx_data"]]
        return read_array_file(kwargs["out_path_res"], t)

    async def verify_result(self, task_result: np.ndarray, **kwargs) -> bool:
        path, t = self.items[kwargs["idx_data"]]
        want = np.sort(read_array_file(path, t), kind="stable")
        return bool(task_result.shape == want.shape and np.array_equal(task_result, want))

    def get_attr(self):
        return {}


LAB_PROCESSORS = {
    "lab1": Lab1Processor,
    "lab2": Lab2Processor,
    "lab3": Lab3Processor,
    "lab5": Lab5Processor,
}
