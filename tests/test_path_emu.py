"""Batched path-group kernel (csrc/path_kernel.cuh + csrc/path_plan.cpp) on the CPU: the kernel SOURCE
runs on 256 host threads (tests/cpp/path_emu.cpp), the factor is compared with the oracle / the
reference's golden vectors.  Covers the planner (paths, groups, position lists, carried blocks, the
dedicated workspace area) and the kernel logic; only the sm_100a build itself needs the GPU tests."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import sparsecholesky_b200 as sc
from oracle import pyoracle as po
from sparsecholesky_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "tests", "cpp", "libpathemu%s.so")
CSRC = os.path.join(ROOT, "sparsecholesky_b200", "csrc")
GOLD = os.path.join(ROOT, "tests", "golden")


def _build(threads, carry_per_thread, opt=0):
    so = SO % (f"{threads}_opt{opt}" if opt else threads)
    srcs = [os.path.join(ROOT, "tests", "cpp", "path_emu.cpp")] + [os.path.join(CSRC, f) for f in ("path_plan.cpp", "plan.cpp", "symbolic.cpp")]
    deps = srcs + [os.path.join(CSRC, f) for f in ("path_plan.h", "path_kernel.cuh", "plan.h", "symbolic.h")]
    if not os.path.exists(so) or any(os.path.getmtime(f) > os.path.getmtime(so) for f in deps):
        subprocess.run(["g++", "-std=c++20", "-O2", "-fPIC", "-shared", "-pthread", f"-DPK_THREADS={threads}",
                        f"-DPK_CARRY_PER_THREAD={carry_per_thread}", f"-DPK_OPT={opt}", "-o", so] + srcs, check=True)
    L = C.CDLL(so)
    L.pathemu_factor.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                 C.c_void_p, C.POINTER(C.c_int), C.c_void_p, C.c_char_p, C.c_int]
    return L


@pytest.fixture(scope="module")
def emu():
    """64 host threads per CTA (a barrier of 256 OS threads costs milliseconds), same register budget for carried blocks."""
    return _build(64, 112)


@pytest.fixture(scope="module")
def emu_opt():
    """every staged compile-time variant switched on (PK_OPT = 15, csrc/path_kernel.cuh)"""
    return _build(64, 112, opt=15)


@pytest.fixture(scope="module")
def emu256():
    """the kernel's real CTA shape: 256 threads, 28 carried entries per thread"""
    return _build(256, 28)


def run(L, A, relax=2, level0=-1, max_nu=160, carry_cap=28 * 256, mode=0):
    S = sc.schol(A, relax=relax)
    Lx = np.empty(S.nnz)
    info = C.c_int(0)
    st = np.zeros(6, np.int64)
    err = C.create_string_buffer(2048)
    rc = L.pathemu_factor(A.n, A.p.ctypes.data, A.i.ctypes.data, A.x.ctypes.data, relax, level0, max_nu, carry_cap, mode,
                          Lx.ctypes.data, C.byref(info), st.ctypes.data, err, 2048)
    assert rc == 0, err.value.decode()
    return Lx, info.value, dict(zip(["level0", "groups", "members", "children", "carried", "max_nu"], (int(v) for v in st)))


def golden(name):
    g = np.load(os.path.join(GOLD, name + ".npz"))
    return sc.CscMatrix(int(g["n"]), g["Ap"], g["Ai"], g["Ax"]), g["Lx"]


def relerr(Lx, ref):
    return float((np.abs(Lx - ref) / np.maximum(np.abs(ref), 1e-300)).max())


def test_1138_bus_batched_setting(emu256):
    # the configuration of BASELINE config 5: relax = 2, the engine's level rule, the kernel's CTA shape
    A, ref = golden("1138_bus")
    Lx, info, st = run(emu256, A)
    assert info == 0 and relerr(Lx, ref) <= 1e-10
    # 174 fronts of the upper levels in 19 groups; the links of the main path are carried in registers
    assert st["members"] == 174 and st["groups"] <= 25 and st["carried"] >= 5 and st["max_nu"] <= 160


@pytest.mark.parametrize("max_nu,carry_cap", [(128, 28 * 256), (236, 28 * 256), (160, 0), (160, 300)])
def test_1138_bus_group_sizes_and_workspace_handoff(emu, max_nu, carry_cap):
    # small groups, one-CTA-per-SM groups, and blocks forced through the dedicated workspace area
    A, ref = golden("1138_bus")
    Lx, info, st = run(emu, A, max_nu=max_nu, carry_cap=carry_cap)
    assert info == 0 and relerr(Lx, ref) <= 1e-10
    if carry_cap == 0:
        assert st["carried"] == 0


@pytest.mark.parametrize("name", ["bcsstk01", "lap2d_24", "lap3d_10", "lap2d_7x5", "test3"])
@pytest.mark.parametrize("relax", [1, 2])
def test_whole_tree_as_path_groups(emu, name, relax):
    # level0 = 0: every front goes through the group program (bushy trees: many side paths, blocks handed over
    # through the workspace); max_nu small enough that long paths are cut
    A, ref = golden(name)
    S = sc.schol(A, relax=relax)
    mm = int(S.panel_layout()["m"].max())
    for max_nu in (mm, mm + 7, 4 * mm):
        groups = []
        for mode in (0, 1):                      # path groups, subtree groups
            Lx, info, st = run(emu, A, relax=relax, level0=0, max_nu=max_nu, mode=mode)
            assert info == 0 and relerr(Lx, ref) <= 1e-10, (max_nu, mode, st)
            groups.append(st["groups"])
        assert groups[1] <= groups[0]            # subtrees never need more groups than paths


def test_1138_bus_subtree_groups(emu):
    A, ref = golden("1138_bus")
    Lx0, info0, st0 = run(emu, A, mode=0)
    Lx1, info1, st1 = run(emu, A, mode=1)
    assert info1 == 0 and relerr(Lx1, ref) <= 1e-10
    assert st1["members"] == 174 and st1["groups"] < st0["groups"]


def test_upper_levels_only_on_a_grid(emu):
    perm = synth.nd_perm(20, 20, 1)
    ti, tj, tx, n = synth.laplacian_triplets(20, 20, 1, perm=perm)
    A = sc.triplet_to_csc_matrix(ti, tj, tx, n)
    rc, Lp, Li, ref = po.chol(A.n, A.p, A.i, A.x)
    assert rc == 0
    for level0 in (0, 1, 2, 3):
        Lx, info, st = run(emu, A, relax=1, level0=level0, max_nu=100)
        assert info == 0 and relerr(Lx, ref) <= 1e-10


def test_not_positive_definite_column(emu):
    A, _ = golden("1138_bus")
    x = A.x.copy()
    j = 900
    x[A.p[j + 1] - 1] = -1.0                      # diagonal of column 900 (rows ascending: the diagonal is last)
    B = sc.CscMatrix(A.n, A.p, A.i, x)
    Lx, info, st = run(emu, B)
    rc, _, _, _ = po.chol(B.n, B.p, B.i, B.x)
    assert rc != 0 and info == j + 1


def test_front_larger_than_the_group_capacity_is_refused(emu):
    A, _ = golden("1138_bus")
    S = sc.schol(A, relax=2)
    err = C.create_string_buffer(512)
    rc = emu.pathemu_factor(A.n, A.p.ctypes.data, A.i.ctypes.data, A.x.ctypes.data, 2, -1, 50, 0, 0, np.empty(S.nnz).ctypes.data,
                            C.byref(C.c_int(0)), None, err, 512)
    assert rc != 0 and b"max_nu" in err.value


@pytest.mark.parametrize("opt", [0, 15])
def test_kernel_source_has_no_data_race_under_thread_sanitizer(tmp_path, opt):
    """Every shared-memory hand-over in the kernel is separated by a barrier: ThreadSanitizer watches the host threads
    that play the CTA (a barrier removed from path_kernel.cuh is reported as a data race here)."""
    tsan = subprocess.run(["g++", "-print-file-name=libtsan.so"], capture_output=True, text=True).stdout.strip()
    if not os.path.isabs(tsan) or not os.path.exists(tsan):
        pytest.skip("libtsan not available")
    so = str(tmp_path / "libpathemu_tsan.so")
    srcs = [os.path.join(ROOT, "tests", "cpp", "path_emu.cpp")] + [os.path.join(CSRC, f) for f in ("path_plan.cpp", "plan.cpp", "symbolic.cpp")]
    subprocess.run(["g++", "-std=c++20", "-O1", "-g", "-fsanitize=thread", "-fPIC", "-shared", "-pthread", "-DPK_THREADS=64",
                    "-DPK_CARRY_PER_THREAD=112", f"-DPK_OPT={opt}", "-o", so] + srcs, check=True)
    code = f"""
import ctypes as C, numpy as np, sys
sys.path.insert(0, {ROOT!r})
g = np.load({os.path.join(GOLD, 'bcsstk01.npz')!r})
L = C.CDLL({so!r})
L.pathemu_factor.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                             C.POINTER(C.c_int), C.c_void_p, C.c_char_p, C.c_int]
Ap, Ai, Ax = (np.ascontiguousarray(g[k]) for k in ("Ap", "Ai", "Ax"))
for max_nu, cap, mode in ((50, 7168, 0), (45, 0, 0), (50, 7168, 1)):      # carried blocks / through the workspace / subtree groups
    Lx = np.empty(g["Lx"].size); info = C.c_int(0); err = C.create_string_buffer(512)
    rc = L.pathemu_factor(int(g["n"]), Ap.ctypes.data, Ai.ctypes.data, Ax.ctypes.data, 1, 0, max_nu, cap, mode, Lx.ctypes.data,
                          C.byref(info), None, err, 512)
    assert rc == 0 and info.value == 0, err.value
    assert float((np.abs(Lx - g["Lx"]) / np.abs(g["Lx"])).max()) <= 1e-10
print("emu ok")
"""
    env = dict(os.environ, LD_PRELOAD=tsan, TSAN_OPTIONS="halt_on_error=0 report_bugs=1 exitcode=0")
    r = subprocess.run([os.sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=600)
    assert "emu ok" in r.stdout, r.stdout + r.stderr[-2000:]
    assert "ThreadSanitizer: data race" not in r.stderr, r.stderr[:3000]



@pytest.mark.parametrize("mode", [0, 1])
def test_staged_kernel_variants(emu_opt, mode):
    # PK_OPT = 15: planner-supplied corner sizes, one warp barrier per pivot, unmasked interior chunks, prefetch (a no-op
    # on the host) -- same factor to rounding on the batched configuration and on whole trees
    A, ref = golden("1138_bus")
    Lx, info, st = run(emu_opt, A, mode=mode)
    assert info == 0 and relerr(Lx, ref) <= 1e-10
    for name in ("bcsstk01", "lap2d_24", "lap3d_10"):
        A, ref = golden(name)
        mm = int(sc.schol(A, relax=1).panel_layout()["m"].max())
        Lx, info, st = run(emu_opt, A, relax=1, level0=0, max_nu=mm + 7, mode=mode)
        assert info == 0 and relerr(Lx, ref) <= 1e-10, (name, st)
