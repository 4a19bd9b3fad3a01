"""A C program that makes the Go shim's call sequence (init -> type flags -> upload -> score one pod ->
delta -> score again) against hand-derived KAT literals: proves the C ABI is usable from plain C99 without
the Python harness, on a real GPU."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "4paradigm-k8s-vgpu-scheduler_b200", "csrc")


def _build(tmp_path):
    exe = str(tmp_path / "driver")
    subprocess.check_call(["gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", "-O1", "-o", exe,
                           os.path.join(ROOT, "tests", "c", "driver.c"), "-I" + os.path.join(ROOT, "include"),
                           "-L" + CSRC, "-lvgpuscore", "-Wl,-rpath," + CSRC])
    return exe


def test_c_driver_compiles_as_c99(built, tmp_path):
    _build(tmp_path)


@pytest.mark.gpu
def test_c_driver_runs_the_filter_sequence(built, tmp_path):
    exe = _build(tmp_path)
    p = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0 and "driver: ok" in p.stdout, p.stdout + p.stderr
