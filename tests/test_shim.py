"""The C++ shim (dynolog_b200/shim/HbtGpuShim.h) keeps the reference's class shapes.
CPU: it compiles against the C ABI.  GPU: a reference-shaped C++ test runs through it."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "shim", "shim_test.cpp")


def _compile(out, link=True):
    cmd = ["g++", "-std=c++17", "-O1", "-Wall", "-Wextra", "-I", os.path.join(ROOT, "include"),
           "-I", os.path.join(ROOT, "dynolog_b200", "shim"), SRC]
    if link:
        lib_dir = os.path.join(ROOT, "dynolog_b200")
        cmd += ["-o", out, "-L", lib_dir, "-lhbtgpu", f"-Wl,-rpath,{lib_dir}"]
    else:
        cmd += ["-c", "-o", out]
    subprocess.check_call(cmd)


def test_shim_compiles(tmp_path):
    _compile(str(tmp_path / "shim_test.o"), link=False)


@pytest.mark.gpu
def test_shim_reference_shaped_test(tmp_path):
    exe = str(tmp_path / "shim_test")
    _compile(exe)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "SHIM OK" in r.stdout, r.stdout + r.stderr
