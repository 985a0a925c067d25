"""Utilities: image-format bridge, downloads, device timing, clock sampling."""
from .imgdata import ImgData, decode_data_bytes, encode_data_bytes, hex_to_bytes, bytes_to_hex
from .download_files import download_file
from .timing import cuda_time_ms, L2Flusher
from .clocks import ClockSampler
from .graphs import GraphedStep

__all__ = [
    "ImgData", "decode_data_bytes", "encode_data_bytes", "hex_to_bytes", "bytes_to_hex",
    "download_file", "cuda_time_ms", "L2Flusher", "ClockSampler", "GraphedStep",
]
