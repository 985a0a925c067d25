"""The harness end to end without a GPU: any executable honouring the stdin/stdout
contract can stand in for the CUDA binary (here: the CPU builds)."""
import asyncio
import os

import pandas as pd
import pytest

from cuda_mpi_openmp_b200.harness import BaseTester, Lab1Processor, Lab2Processor, Lab3Processor, Lab5Processor


def run(tester, proc):
    return asyncio.run(tester.run_experiments(lab_processor=proc))


@pytest.fixture()
def in_repo(built, monkeypatch):
    monkeypatch.chdir(built)
    return built


def test_lab2_harness_checks_goldens(in_repo, tmp_path):
    exe = str(in_repo / "lab2/src/cpu_exe")
    proc = Lab2Processor(dir_to_data="./lab2/data", dir_to_data_out=str(tmp_path / "out"))
    assert len(proc.data_output_gt) == 2                      # test_01 / test_02 have goldens
    n = len(proc.data_input)
    tester = BaseTester(exe, k_times=n, kernel_sizes=[[None, None]], metadata_columns2plot=["filename"])
    df = run(tester, proc)
    assert df.shape[0] == n and df["test_verification_result"].all() and (df["device"] == "CUDA").all()
    stats = pd.read_csv(in_repo / "lab2/src/stats_cpu_exe.csv")
    assert {"idx_run_time", "bin_name", "kernel_size", "time_kernel_exe_ms", "status", "filename"} <= set(stats.columns)
    assert os.path.exists(in_repo / "lab2/src/median_execution_time.png")


def test_lab2_harness_detects_wrong_output(in_repo, tmp_path):
    # hw2 reads junk and exits 0 without writing the output image -> every row fails
    proc = Lab2Processor(dir_to_data="./lab2/data", dir_to_data_out=str(tmp_path / "out"))
    tester = BaseTester(str(in_repo / "lab3/src/test_read_input"), k_times=1, kernel_sizes=[[None, None]],
                        metadata_columns2plot=[])
    df = run(tester, proc)
    assert df.shape[0] == 0 and os.path.exists(in_repo / "lab3/src/failed_test_read_input.csv")
    os.remove(in_repo / "lab3/src/failed_test_read_input.csv")


def test_lab3_harness(in_repo, tmp_path):
    proc = Lab3Processor(dir_to_data="./lab3/data", dir_to_data_out=str(tmp_path / "out"))
    assert len(proc.data_output_gt) == 1
    tester = BaseTester(str(in_repo / "lab3/src/cpu_exe"), k_times=len(proc.data_input),
                        kernel_sizes=[[None, None]], metadata_columns2plot=["filename"])
    df = run(tester, proc)
    assert df["test_verification_result"].all()


def test_lab1_harness_with_cpu_comparator(in_repo):
    exe = str(in_repo / "lab1/src/cpu_exe_to_plot")
    tester = BaseTester(exe, k_times=2, kernel_sizes=[[None, None]], metadata_columns2plot=["vector_size"],
                        binary_path_cpu=exe)
    df = run(tester, Lab1Processor(min_vector_size=1500, max_vector_size=1600))
    assert df.shape[0] == 4 and set(df["device"]) == {"CUDA", "CPU"} and df["test_verification_result"].all()


def test_lab5_harness(in_repo, tmp_path):
    proc = Lab5Processor(dir_to_data="./lab5/data", dir_to_data_out=str(tmp_path / "out"), extra_random=1,
                         random_size=5000)
    tester = BaseTester(str(in_repo / "lab5/src/to_plot_cpu_exe"), k_times=len(proc.items),
                        kernel_sizes=[[None, None]], metadata_columns2plot=["filename"])
    df = run(tester, proc)
    assert df.shape[0] == 6 and df["test_verification_result"].all()


def test_kernel_size_prefix_encoding(in_repo, tmp_path):
    # list sizes become "a\nb" and are prepended (ref: tester.py:113-121)
    class Echo(Lab1Processor):
        async def pre_process(self, **kw):
            return "2\n1 2\n3 4", {"first_vector_printed": None, "second_vector_printed": None}, {}

        async def verify_result(self, task_result, **kw):
            return True

    tester = BaseTester(str(in_repo / "lab1/src/cpu_exe_to_plot"), k_times=1, kernel_sizes=[[[32, 32], [16, 16]]],
                        metadata_columns2plot=[], return_inp=True)
    proc = Echo()
    from cuda_mpi_openmp_b200.harness import run_subprocess
    res = asyncio.run(run_subprocess("/bin/cat", proc, True, [32, 32], [16, 16]))
    assert res.debug_data["input_str"].startswith("32\n32\n16\n16\n2\n")
