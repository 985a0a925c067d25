"""Fetch an extra test image over HTTP (ref: utils/download_files.py:5-35).

Only used for the processors' ``extra_links_to_png`` option; there is no network
on the build/GPU boxes, so failures surface as a normal exception.
"""
from __future__ import annotations

import os
import shutil
from urllib.parse import urlparse


def download_file(url: str, save_dir: str, filename: str | None = None, timeout: float = 30.0) -> str:
    os.makedirs(save_dir, exist_ok=True)
    name = filename or os.path.basename(urlparse(url).path) or "download.bin"
    dest = os.path.join(save_dir, name)
    if url.startswith("file://") or os.path.exists(url):  # local path: handy offline
        shutil.copyfile(url[7:] if url.startswith("file://") else url, dest)
        return dest
    import requests

    with requests.get(url, stream=True, timeout=timeout) as resp:
        resp.raise_for_status()
        tmp = dest + ".part"
        with open(tmp, "wb") as f:
            for chunk in resp.iter_content(chunk_size=1 << 16):
                if chunk:
                    f.write(chunk)
        os.replace(tmp, dest)
    return dest
