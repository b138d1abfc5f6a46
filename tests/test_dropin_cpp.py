"""The C++23 drop-in header (include/chol.hpp, include/mtx_reader.hpp) compiled against libspchol.so:
the reference's own gtest cases restated as plain asserts (tests/cpp/test_dropin.cpp)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tests", "cpp", "test_dropin.bin")


def _build():
    from sparsecholesky_b200 import build
    build.build()
    libdir = os.path.join(ROOT, "sparsecholesky_b200")
    cmd = ["g++", "-std=c++23", "-O1", "-Wall", "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "cpp", "test_dropin.cpp"), "-o", EXE, "-L", libdir, "-lspchol",
           f"-Wl,-rpath,{libdir}"]
    subprocess.run(cmd, check=True, capture_output=True, text=True)


def test_dropin_header_host_side():
    _build()
    r = subprocess.run([EXE, os.path.join(ROOT, "tests", "data")], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "test_dropin OK" in r.stdout


@pytest.mark.gpu
def test_dropin_header_gpu():
    _build()
    r = subprocess.run([EXE, os.path.join(ROOT, "tests", "data"), "gpu"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "test_dropin OK (gpu)" in r.stdout
