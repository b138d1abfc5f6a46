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


def test_reference_helpers_bit_for_bit(tmp_path):
    """random_sparse (both overloads), the csc_matrix(A, S) constructor and schol_etree of the drop-in header against what
    the reference's own header produced (tests/golden/boundary.npz, generated through oracle/_ref/libref.so)."""
    import numpy as np
    _build()
    out = tmp_path / "dump.txt"
    r = subprocess.run([EXE, os.path.join(ROOT, "tests", "data"), "dump", str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    g = np.load(os.path.join(ROOT, "tests", "golden", "boundary.npz"))
    got = {}
    for line in open(out):
        name, cnt, *vals = line.split()
        got[name] = np.array([float(v) for v in vals])
        assert len(vals) == int(cnt)
    assert set(got) == set(g.files)
    for k in g.files:
        assert np.array_equal(got[k], g[k].astype(np.float64)), k          # bit for bit (same generator, same distribution code)


REF_TESTS = "/root/reference/tests/test_chol.cpp"
REF_EXE = os.path.join(ROOT, "tests", "cpp", "ref_tests.bin")


def build_ref_tests():
    """The reference's tests/test_chol.cpp compiled VERBATIM against include/chol.hpp (GoogleTest macros from the stand-in
    tests/cpp/gtest_shim, dpotrf_ from the dense stand-in in tests/cpp/ref_tests_main.cpp).  Only where the reference tree
    exists (the build container); the binary travels to the GPU box like the other built artefacts."""
    from sparsecholesky_b200 import build
    build.build()
    libdir = os.path.join(ROOT, "sparsecholesky_b200")
    cmd = ["g++", "-std=c++23", "-O1", "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "tests", "cpp", "gtest_shim"),
           REF_TESTS, os.path.join(ROOT, "tests", "cpp", "ref_tests_main.cpp"), "-o", REF_EXE, "-L", libdir, "-lspchol",
           f"-Wl,-rpath,{libdir}"]
    subprocess.run(cmd, check=True, capture_output=True, text=True)


def test_reference_tests_compile_verbatim_and_host_cases_pass():
    if not os.path.exists(REF_TESTS):
        pytest.skip("reference tree absent (GPU box): the binary was built in the build container")
    build_ref_tests()
    r = subprocess.run([REF_EXE, "host"], capture_output=True, text=True)
    assert r.returncode == 0 and "2 ran, 0 failed" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
def test_reference_tests_all_four_pass_on_gpu():
    """EliminationTree, ColumnReach, SimplicialCholesky and SupernodalCholesky (the last one FAILS in the reference itself,
    SURVEY.md section 4: its chol_sn is unfinished; here chol_sn is the GPU supernodal factorization)."""
    if os.path.exists(REF_TESTS):
        build_ref_tests()
    if not os.path.exists(REF_EXE):
        pytest.skip("tests/cpp/ref_tests.bin was not built (needs /root/reference at build time)")
    r = subprocess.run([REF_EXE], capture_output=True, text=True)
    assert r.returncode == 0 and "4 ran, 0 failed" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
def test_dropin_header_gpu():
    _build()
    r = subprocess.run([EXE, os.path.join(ROOT, "tests", "data"), "gpu"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "test_dropin OK (gpu)" in r.stdout
