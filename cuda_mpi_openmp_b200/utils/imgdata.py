"""The lab image container in its three interchangeable encodings.

* ``.data`` — ``int32 w, int32 h`` then ``w*h`` RGBA bytes, row-major (what the binaries read);
* ``.txt``  — the same bytes as hex, 8 hex digits per word, whitespace-insensitive;
* ``.png``  — ordinary image (alpha forced to 255 when converting *from* PNG).

Behavioural parity with ref: utils/converter.py:17-148 (same attribute names so
processors written against it keep working), but the conversions are vectorised
with numpy instead of per-pixel Python loops, and ``out_dir`` lets callers keep
the sibling files out of a read-only or version-controlled input directory.
"""
from __future__ import annotations

import binascii
import os
import struct
import sys
from typing import Optional

import numpy as np


def decode_data_bytes(blob: bytes) -> np.ndarray:
    """``.data`` bytes -> uint8 array [h, w, 4]."""
    if len(blob) < 8:
        raise ValueError("image blob is shorter than its 8-byte header")
    w, h = struct.unpack_from("<ii", blob, 0)
    if w <= 0 or h <= 0 or len(blob) < 8 + 4 * w * h:
        raise ValueError(f"bad image blob: header says {w}x{h}, {len(blob) - 8} pixel bytes present")
    return np.frombuffer(blob, dtype=np.uint8, count=4 * w * h, offset=8).reshape(h, w, 4)


def encode_data_bytes(rgba: np.ndarray) -> bytes:
    """uint8 [h, w, 4] -> ``.data`` bytes."""
    rgba = np.ascontiguousarray(rgba, dtype=np.uint8)
    h, w, c = rgba.shape
    if c != 4:
        raise ValueError("expected an RGBA array")
    return struct.pack("<ii", w, h) + rgba.tobytes()


def hex_to_bytes(text: str) -> bytes:
    return binascii.unhexlify("".join(text.split()))


def bytes_to_hex(blob: bytes) -> str:
    """8 hex digits per group, single spaces, no newlines (ref: utils/converter.py:131-139)."""
    h = binascii.hexlify(blob).decode("ascii")
    return " ".join(h[i:i + 8] for i in range(0, len(h), 8))


class ImgData:
    """Loads one image from any of the three encodings and materialises the others."""

    def __init__(self, path2data: str, idx: Optional[int] = None, out_dir: Optional[str] = None,
                 materialize: bool = True):
        if not os.path.exists(path2data):
            raise AssertionError(f"os.path.exists({path2data}) is False")
        self.idx = idx
        self.dir2save = out_dir if out_dir is not None else os.path.dirname(path2data)
        base = os.path.basename(path2data)
        self.data_name, self.data_ext = os.path.splitext(base)
        self.png = None
        self.hex: Optional[str] = None
        self.c_data_bytes_path = self.hex_path = self.png_path = None
        if out_dir:
            os.makedirs(out_dir, exist_ok=True)

        if self.data_ext == ".data":
            with open(path2data, "rb") as f:
                self.c_data_bytes = f.read()
            self.c_data_bytes_path = path2data
            self.hex = bytes_to_hex(self.c_data_bytes)
            if materialize:
                self.hex_path = self._write_text(".txt", self.hex)
                self.png_path = self._write_png()
        elif self.data_ext == ".png":
            self.c_data_bytes = self._read_png(path2data)
            self.png_path = path2data
            self.hex = bytes_to_hex(self.c_data_bytes)
            # the .data file is always needed: it is what the binaries consume
            self.c_data_bytes_path = self._write_bytes(".data", self.c_data_bytes)
            if materialize:
                self.hex_path = self._write_text(".txt", self.hex)
        elif self.data_ext == ".txt":
            with open(path2data, "r") as f:
                raw = f.read()
            self.hex = raw.replace("\n", " ")
            self.c_data_bytes = hex_to_bytes(raw)
            self.hex_path = path2data
            self.c_data_bytes_path = self._write_bytes(".data", self.c_data_bytes)
            if materialize:
                self.png_path = self._write_png()
        else:
            raise ValueError(
                "`path2data` should endwith '.data' OR '.png' OR '.txt'. \n"
                f"NOW: `path2data`={path2data}")
        # KB including the bytes-object overhead, as the reference reports it
        # (ref: utils/converter.py:11-13) so filename labels stay comparable
        self.size: float = sys.getsizeof(self.c_data_bytes) / 1024

    # -- views -------------------------------------------------------------
    @property
    def rgba(self) -> np.ndarray:
        return decode_data_bytes(self.c_data_bytes)

    @property
    def shape(self) -> tuple[int, int]:
        w, h = struct.unpack_from("<ii", self.c_data_bytes, 0)
        return h, w

    # -- writers -----------------------------------------------------------
    def _path(self, ext: str) -> str:
        return os.path.join(self.dir2save, f"{self.data_name}{ext}")

    def _write_bytes(self, ext: str, blob: bytes) -> str:
        p = self._path(ext)
        with open(p, "wb") as f:
            f.write(blob)
        return p

    def _write_text(self, ext: str, text: str) -> str:
        p = self._path(ext)
        with open(p, "w") as f:
            f.write(text)
        return p

    def _write_png(self) -> Optional[str]:
        try:
            from PIL import Image
        except ImportError:  # PNG view is a convenience, not part of the contract
            return None
        self.png = Image.fromarray(self.rgba, mode="RGBA")
        p = self._path(".png")
        self.png.save(p)
        return p

    def _read_png(self, path: str) -> bytes:
        from PIL import Image

        img = Image.open(path)
        self.png = img
        rgb = np.asarray(img.convert("RGB"), dtype=np.uint8)
        rgba = np.empty(rgb.shape[:2] + (4,), dtype=np.uint8)
        rgba[..., :3] = rgb
        rgba[..., 3] = 255
        return encode_data_bytes(rgba)
