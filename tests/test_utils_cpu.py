import os

import numpy as np

from cuda_mpi_openmp_b200.harness import parse_kernel_time, unknown_args_to_kwargs
from cuda_mpi_openmp_b200.utils import ImgData, bytes_to_hex, decode_data_bytes, encode_data_bytes, hex_to_bytes


def test_time_line_regex():
    assert parse_kernel_time("CUDA execution time: <0.134560 ms>") == 0.13456
    assert parse_kernel_time("CPU execution time: <8.3 ms>\nrest") == 8.3
    assert parse_kernel_time("FINISHED!") is None


def test_unknown_args():
    kw = unknown_args_to_kwargs(["--min_vector_size", "10", "--flag", "--ratio", "0.5", "--name", "abc",
                                 "--on", "True", "stray", "--dir_to_data", "./lab2/data"])
    assert kw == {"min_vector_size": 10, "flag": True, "ratio": 0.5, "name": "abc", "on": True,
                  "dir_to_data": "./lab2/data"}


def test_hex_roundtrip():
    blob = encode_data_bytes(np.arange(24, dtype=np.uint8).reshape(2, 3, 4))
    text = bytes_to_hex(blob)
    assert text.startswith("03000000 02000000 00010203")
    assert hex_to_bytes(text + "\n") == blob
    assert np.array_equal(decode_data_bytes(blob), np.arange(24, dtype=np.uint8).reshape(2, 3, 4))


def test_imgdata_three_encodings(repo_root, tmp_path):
    a = ImgData(str(repo_root / "lab2/data/test_01.txt"), out_dir=str(tmp_path))
    assert a.shape == (3, 3) and os.path.exists(a.c_data_bytes_path) and os.path.exists(a.png_path)
    b = ImgData(a.c_data_bytes_path, out_dir=str(tmp_path / "b"))
    assert b.c_data_bytes == a.c_data_bytes and "".join(b.hex.split()) == "".join(a.hex.split())
    c = ImgData(b.png_path, out_dir=str(tmp_path / "c"))
    assert c.rgba[..., :3].tolist() == a.rgba[..., :3].tolist() and (c.rgba[..., 3] == 255).all()
    assert abs(a.size - 0.07520) < 1e-4     # same KB figure the reference prints for a 44-byte file
    assert not any(f.endswith(".png") for f in os.listdir(repo_root / "lab2/data") if f.startswith("test_01"))


def test_png_conversion_matches_fixture(repo_root, tmp_path):
    png = ImgData(str(repo_root / "lab2/data/lenna.png"), out_dir=str(tmp_path), materialize=False)
    with open(repo_root / "lab2/test_data/lenna.data", "rb") as f:
        assert png.c_data_bytes == f.read()


def test_lab3_sample_points_outside_the_image_are_an_error():
    """A point outside the whole image would be skipped by every band's kernel and silently
    drop out of the class statistics: pack_points refuses it on the host."""
    import pytest
    from cuda_mpi_openmp_b200.ops.classify import pack_points

    ok = [np.array([[0, 0], [9, 4]]), np.array([[3, 2]])]
    pts, off = pack_points(ok, w=10, h=5)
    assert pts.tolist() == [0, 0, 9, 4, 3, 2] and off.tolist() == [0, 2, 3]
    for bad in ([[10, 0]], [[0, 5]], [[-1, 2]], [[2, -1]]):
        with pytest.raises(ValueError, match="outside"):
            pack_points([np.array([[1, 1]]), np.array(bad)], w=10, h=5)
    pack_points([np.array([[10, 0]])])          # no bounds given: packing only
