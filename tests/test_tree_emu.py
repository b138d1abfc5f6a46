"""Warp-per-subtree kernel (csrc/tree_kernel.cuh + csrc/tree_plan.cpp) on the CPU: the kernel SOURCE runs on 32 host
threads per warp (tests/cpp/tree_emu.cpp), the factor is compared with the oracle / the reference's golden vectors.
Covers the planner (postorder tasks, carried blocks, stash stack, reversed maps, A entry lists) and the kernel logic
(in-place spreading of a carried block, pivots, rank-w update, the three exits of a contribution block); only the
sm_100a build itself needs the GPU tests."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import sparsecholesky_b200 as sc
from oracle import pyoracle as po
from sparsecholesky_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "sparsecholesky_b200", "csrc")
GOLD = os.path.join(ROOT, "tests", "golden")


def _build(tsan=False):
    so = os.path.join(ROOT, "tests", "cpp", "libtreeemu_tsan.so" if tsan else "libtreeemu.so")
    srcs = [os.path.join(ROOT, "tests", "cpp", "tree_emu.cpp")] + [os.path.join(CSRC, f) for f in ("tree_plan.cpp", "plan.cpp", "symbolic.cpp")]
    deps = srcs + [os.path.join(CSRC, f) for f in ("tree_plan.h", "tree_kernel.cuh", "plan.h", "symbolic.h")]
    if not os.path.exists(so) or any(os.path.getmtime(f) > os.path.getmtime(so) for f in deps):
        flags = ["-fsanitize=thread", "-O1", "-g"] if tsan else ["-O2"]
        subprocess.run(["g++", "-std=c++20", "-fPIC", "-shared", "-pthread", "-o", so] + flags + srcs, check=True)
    return so


@pytest.fixture(scope="module")
def emu():
    L = C.CDLL(_build())
    L.treeemu_factor.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int,
                                 C.c_void_p, C.POINTER(C.c_int), C.c_void_p, C.c_char_p, C.c_int]
    return L


def run(L, A, whole, relax=1, cap_m=117, work=2.5e6, nw=1):
    S = sc.schol(A, relax=relax)
    Lx = np.empty(S.nnz)
    info = C.c_int(0)
    st = np.zeros(8, np.int64)
    err = C.create_string_buffer(2048)
    rc = L.treeemu_factor(A.n, A.p.ctypes.data, A.i.ctypes.data, A.x.ctypes.data, relax, int(whole), cap_m, work, nw,
                          Lx.ctypes.data, C.byref(info), st.ctypes.data, err, 2048)
    assert rc == 0, err.value.decode()
    return Lx, info.value, dict(zip(["tasks", "fronts", "stashes", "stash_doubles", "cap_m", "ns", "groups", "padded_flops"], (int(v) for v in st)))


def golden(name):
    g = np.load(os.path.join(GOLD, name + ".npz"))
    return sc.CscMatrix(int(g["n"]), g["Ap"], g["Ai"], g["Ax"]), g["Lx"]


def relerr(Lx, ref):
    return float((np.abs(Lx - ref) / np.maximum(np.abs(ref), 1e-300)).max())


@pytest.mark.parametrize("name", ["test3", "bcsstk01", "1138_bus", "lap2d_24", "lap2d_7x5"])      # (lap3d_10 has fronts beyond 128 rows)
@pytest.mark.parametrize("relax,nw", [(0, 1), (1, 4), (1, 2)])
def test_whole_tree_vs_reference_golden(emu, name, relax, nw):
    # batched mode: one team walks the whole tree and writes the CSC factor itself
    A, ref = golden(name)
    Lx, info, st = run(emu, A, whole=True, relax=relax, nw=nw)
    assert info == 0 and relerr(Lx, ref) <= 1e-10
    assert st["tasks"] == 1 and st["fronts"] == st["ns"]


@pytest.mark.parametrize("name", ["test3", "bcsstk01", "1138_bus", "lap2d_24", "lap2d_7x5"])
@pytest.mark.parametrize("cap_m,nw", [(117, 4), (128, 1), (64, 2), (0, 4)])
def test_group_program_vs_reference_golden(emu, name, cap_m, nw):
    # the batched program over amalgamated groups: fewer, bigger dense steps with explicit zeros that must stay exact zeros
    A, ref = golden(name)
    cap = max(cap_m, 1)                              # (cap 1: nothing is absorbed, every front its own group)
    Lx, info, st = run(emu, A, whole=2, cap_m=cap, nw=nw)
    assert info == 0 and relerr(Lx, ref) <= 1e-10
    assert st["tasks"] == 1 and 1 <= st["groups"] <= st["ns"] and st["fronts"] == st["groups"]
    if name == "1138_bus" and cap_m == 117:
        assert st["groups"] <= 60                    # 529 fronts -> ~52 groups


@pytest.mark.parametrize("name", ["bcsstk01", "1138_bus", "lap2d_24", "lap3d_10"])
@pytest.mark.parametrize("cap_m,work,nw", [(117, 2.5e6, 1), (117, 2.5e6, 4), (40, 2e4, 2), (16, 1e3, 1)])
def test_subtrees_vs_reference_golden(emu, name, cap_m, work, nw):
    # single-matrix mode: subtree tasks hand their root blocks to the level program in its own layout
    A, ref = golden(name)
    Lx, info, st = run(emu, A, whole=False, cap_m=cap_m, work=work, nw=nw)
    assert info == 0 and relerr(Lx, ref) <= 1e-10
    assert st["fronts"] <= st["ns"]


@pytest.mark.parametrize("name", ["bcsstk01", "1138_bus", "lap2d_24", "lap3d_10"])
@pytest.mark.parametrize("cap1,work1,nw1,tier2", [(16, 1e3, 1, "117,3e7,4"), (24, 2e4, 1, "64,1e5,2"), (8, 1e3, 2, "40,2e4,1")])
def test_two_tiers_vs_reference_golden(emu, name, cap1, work1, nw1, tier2, monkeypatch):
    # a second tier of subtree tasks on top of the first: its leaves read the first tier's root blocks from the level
    # program's workspace (column-major), everything above it stays with the level program
    monkeypatch.setenv("TREEEMU_TIER2", tier2)
    A, ref = golden(name)
    Lx, info, st = run(emu, A, whole=False, cap_m=cap1, work=work1, nw=nw1)
    assert info == 0 and relerr(Lx, ref) <= 1e-10
    if tier2.startswith("117"):
        assert st["groups"] >= 1 and st["padded_flops"] >= 1      # (tasks, fronts) of the second tier
    assert st["fronts"] + st["padded_flops"] <= st["ns"]


@pytest.mark.parametrize("team", [1, 2, 3])
def test_serial_warp_rotation(emu, team, monkeypatch):
    # on the GPU the warp that runs the serial 8 x 8 factorizations is blockIdx.x mod NW; the host run picks it by hand
    monkeypatch.setenv("TREEEMU_TEAM", str(team))
    A, ref = golden("1138_bus")
    Lx, info, st = run(emu, A, whole=2, cap_m=117, nw=4)
    assert info == 0 and relerr(Lx, ref) <= 1e-10
    A, ref = golden("lap2d_24")
    Lx, info, st = run(emu, A, whole=False, cap_m=40, work=2e4, nw=2)
    assert info == 0 and relerr(Lx, ref) <= 1e-10


def test_subtrees_on_a_larger_grid(emu):
    dims = (48, 40, 1)
    ti, tj, tx, n = synth.laplacian_triplets(*dims, perm=synth.nd_perm(*dims))
    A = sc.triplet_to_csc_matrix(ti, tj, tx, n)
    rc, _, _, ref = po.chol(n, A.p, A.i, A.x)
    assert rc == 0
    for cap_m, work, nw in ((117, 2e5, 4), (64, 2e4, 1)):
        Lx, info, st = run(emu, A, whole=False, cap_m=cap_m, work=work, nw=nw)
        assert info == 0 and relerr(Lx, ref) <= 1e-10 and st["tasks"] >= 2 and st["stashes"] > 0
    os.environ["TREEEMU_TIER2"] = "64,2e5,4"
    try:
        Lx, info, st = run(emu, A, whole=False, cap_m=32, work=2e4, nw=1)
    finally:
        del os.environ["TREEEMU_TIER2"]
    assert info == 0 and relerr(Lx, ref) <= 1e-10 and st["groups"] >= 2


def test_not_positive_definite_column_is_reported(emu):
    A, _ = golden("bcsstk01")
    x = A.x.copy()
    j = 20
    x[A.p[j + 1] - 1] = -1.0                    # diagonal of column j (last entry of an upper-CSC column)
    B = sc.CscMatrix(A.n, A.p, A.i, x)
    rc, _, _, _ = po.chol(B.n, B.p, B.i, B.x)
    _, info, _ = run(emu, B, whole=True, nw=4)
    assert rc > 0 and info == rc               # same 1-based column as the oracle (reference test chol.hpp:849)
    _, info, _ = run(emu, B, whole=2, nw=4)    # ... also through the column table of an amalgamated group
    assert info == rc


def test_thread_sanitizer_finds_no_race():
    """the same source under -fsanitize=thread: a missing __syncwarp is a data race between lanes"""
    so = _build(tsan=True)
    code = (
        "import ctypes as C, numpy as np, os, sys\n"
        f"sys.path.insert(0, {ROOT!r})\n"
        "import sparsecholesky_b200 as sc\n"
        f"g = np.load(os.path.join({GOLD!r}, 'bcsstk01.npz'))\n"
        "A = sc.CscMatrix(int(g['n']), g['Ap'], g['Ai'], g['Ax'])\n"
        f"L = C.CDLL({so!r})\n"
        "L.treeemu_factor.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int, C.c_void_p, C.POINTER(C.c_int), C.c_void_p, C.c_char_p, C.c_int]\n"
        "for whole, cap, work, nw in ((2, 117, 2.5e6, 4), (1, 117, 2.5e6, 4), (0, 16, 1e3, 2), (1, 117, 2.5e6, 1)):\n"
        "    Lx = np.empty(len(g['Lx'])); info = C.c_int(0); st = np.zeros(8, np.int64); err = C.create_string_buffer(512)\n"
        "    rc = L.treeemu_factor(A.n, A.p.ctypes.data, A.i.ctypes.data, A.x.ctypes.data, 1, whole, cap, work, nw, Lx.ctypes.data, C.byref(info), st.ctypes.data, err, 512)\n"
        "    assert rc == 0, err.value\n"
        "    assert np.abs(Lx - g['Lx']).max() <= 1e-10 * np.abs(g['Lx']).max()\n"
        "print('TSAN_RUN_OK')\n")
    import glob
    tsan = sorted(glob.glob("/usr/lib/gcc/x86_64-linux-gnu/*/libtsan.so")) + sorted(glob.glob("/usr/lib/x86_64-linux-gnu/libtsan.so*"))
    if not tsan:
        pytest.skip("libtsan not installed")
    env = dict(os.environ, LD_PRELOAD=tsan[0], TREEEMU_TIER2="117,3e7,4", TSAN_OPTIONS="halt_on_error=0 report_signal_unsafe=0 exitcode=66")
    r = subprocess.run(["python", "-c", code], capture_output=True, text=True, env=env, timeout=600)
    assert "TSAN_RUN_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]
    assert "data race" not in r.stderr, r.stderr[-6000:]
