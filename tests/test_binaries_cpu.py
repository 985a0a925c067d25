"""Contract checks of the CPU binaries and utilities (stdin/stdout formats)."""
import struct
import subprocess

import numpy as np
import pytest


def run(exe, stdin, binary=False):
    r = subprocess.run([str(exe)], input=stdin, capture_output=True, text=not binary, timeout=60)
    assert r.returncode == 0, r.stderr
    return r.stdout


@pytest.mark.parametrize("stdin,want", [
    ("1 -3 2", "2.000000 1.000000\n"), ("0 0 0", "any\n"), ("0 0 5", "incorrect\n"),
    ("0 2 1", "-0.500000\n"), ("1 2 1", "-1.000000\n"), ("1 0 1", "imaginary\n"),
])
def test_hw1(built, stdin, want):
    assert run(built / "hw1/src/hw1", stdin) == want


def test_hw2(built):
    assert run(built / "hw2/src/hw2", "5\n3 1 2 5 4") == "1.000000e+00 2.000000e+00 3.000000e+00 4.000000e+00 5.000000e+00 \n"
    assert run(built / "hw2/src/hw2", "0\n") == "\n"


def test_lab1_cpu_contract(built):
    out = run(built / "lab1/src/cpu_exe_to_plot", "3\n1 2 3.5\n0.5 1e100 -2")
    first, rest = out.split("\n", 1)
    assert first.startswith("CPU execution time: <") and first.endswith(" ms>")
    assert rest == "5.0000000000e-01 -1.0000000000e+100 5.5000000000e+00 "
    assert run(built / "lab1/src/cpu_exe", "2\n1 2\n3 5") == "-2.0000000000e+00 -3.0000000000e+00 "


def test_lab3_read_input(built):
    out = run(built / "lab3/src/test_read_input", "2\n2 1 2 3 4\n1 9 8")
    assert "class 1: 2 points" in out and "(3, 4)" in out and "class 2: 1 points" in out and "(9, 8)" in out


@pytest.mark.parametrize("name,exe,fmt", [("int10", "int", "<10i"), ("float10", "float", "<10f"), ("uchar10", "uchar", "<10B")])
def test_lab5_cpu_fixture(built, name, exe, fmt):
    blob = (built / "lab5/data" / name).read_bytes()
    out = run(built / f"lab5/src/cpu_exe_{exe}", blob, binary=True)
    vals = struct.unpack(fmt, out)
    keys = struct.unpack(fmt, blob[4:])
    assert list(vals) == sorted(keys)


def test_lab2_cpu_error_on_missing_file(built):
    r = subprocess.run([str(built / "lab2/src/cpu_exe")], input="/nonexistent.data\n/tmp/x.data", capture_output=True, text=True)
    assert r.returncode != 0 and "cannot open" in r.stderr
