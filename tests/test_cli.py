"""The native benchmark CLI (bin/bench_reduce): flag handling on CPU, full runs on GPU.
Reference counterpart: reduction.cpp:84-204 (main), :31-41 (documented flags), shrQATest.h (PASSED/FAILED/WAIVED protocol)."""
import os
import re
import subprocess

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(REPO, "bin", "bench_reduce")


def run(*args, cwd=None):
    if not os.path.exists(EXE):
        subprocess.run(["make", "-C", REPO, "-j8", "lib", "bins"], check=True, capture_output=True)
    return subprocess.run([EXE, *args], capture_output=True, text=True, timeout=300, cwd=cwd or REPO)


def test_missing_method_is_an_error_like_the_reference():
    r = run("--type=int", "--nolog")
    assert r.returncode == 1 and "MISSING --method FLAG" in r.stderr  # reduction.cpp:124-128


def test_arguments_must_start_with_a_dash():
    r = run("method=SUM")
    assert r.returncode == 1 and "Invalid command line argument" in r.stderr  # cmd_arg_reader.cpp:44-64


def test_unknown_method_and_type_are_rejected():
    assert run("--method=AVG", "--nolog").returncode == 1
    assert run("--method=SUM", "--type=half", "--nolog").returncode == 1


def test_no_gpu_is_waived_not_failed():
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = run("--method=SUM", "--nolog")
    assert r.returncode == 0 and "WAIVED" in r.stdout  # shrQATest.h QA_WAIVED -> exit code 0


@pytest.mark.gpu
@pytest.mark.parametrize("method", ["SUM", "min", "Max"])  # case-insensitive here (the reference's strcmp is not)
@pytest.mark.parametrize("typ", ["int", "float", "double", "bf16"])
def test_reference_output_lines(method, typ, tmp_path):
    r = run(f"--method={method}", f"--type={typ}", "--n=1048576", "--iters=5", cwd=str(tmp_path))
    assert r.returncode == 0, r.stdout + r.stderr
    assert re.search(r"Reduction, Throughput = [0-9.]+ GB/s, Time = [0-9.]+ s, Size = 1048576 Elements, NumDevsUsed = 1, Workgroup = \d+", r.stdout)
    assert "GPU result =" in r.stdout and "CPU result =" in r.stdout and r.stdout.strip().endswith("PASSED")
    assert os.path.exists(tmp_path / "reduction.txt") and os.path.exists(tmp_path / "SdkMasterLog.csv")  # shrLog / MASTER sinks
    assert "Throughput" in open(tmp_path / "SdkMasterLog.csv").read()


@pytest.mark.gpu
def test_non_power_of_two_cpufinal_kernels_and_shmoo(tmp_path):
    # the reference only supports powers of two, reads out of bounds otherwise for min/max, adds in its --cpufinal
    # fall-back even for min/max, launches nothing for --kernel=0..5 and aborts on --shmoo
    for extra in (["--cpufinal"], ["--cputhresh=7"], ["--kernel=1"], ["--kernel=2"], ["--kernel=3", "--cluster=2"], []):
        for method in ("MIN", "MAX", "SUM"):
            r = run(f"--method={method}", "--type=double", "--n=1000003", "--iters=3", "--nolog", "--fill=signed", *extra)
            assert r.returncode == 0 and "PASSED" in r.stdout, (extra, method, r.stdout[-400:])
    r = run("--method=SUM", "--type=float", "--shmoo", "--n=1048576", "--iters=3", "--nolog", f"--csv={tmp_path}/s.csv")
    assert r.returncode == 0 and "PASSED" in r.stdout and len(open(tmp_path / "s.csv").read().splitlines()) >= 4 * 6


@pytest.mark.gpu
def test_qatest_protocol():
    r = run("--method=SUM", "--n=65536", "--iters=2", "--nolog", "--qatest")
    assert r.returncode == 0 and "&&&& RUNNING" in r.stderr and "&&&& PASSED" in r.stderr  # shrQATest.h:96-109,143-186
