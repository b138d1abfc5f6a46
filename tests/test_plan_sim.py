"""CPU check of the numeric launch program, single- and multi-rank: the product's planner
(csrc/plan.cpp) builds the plan of every rank, tests/cpp/plan_sim.cpp interprets the ops on the host
(transfers between simulated ranks included) and the resulting L is compared with the oracle.
This is how the multi-GPU host logic (front mapping, block-column ownership, transfer order,
workspace lifetimes) is covered without a GPU; the CUDA kernels themselves are covered by -m gpu."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import sparsecholesky_b200 as sc
from oracle import pyoracle as po
from sparsecholesky_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "tests", "cpp", "libplansim.so")
CSRC = os.path.join(ROOT, "sparsecholesky_b200", "csrc")
GOLD = os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="module")
def sim():
    srcs = [os.path.join(ROOT, "tests", "cpp", "plan_sim.cpp"), os.path.join(CSRC, "plan.cpp"), os.path.join(CSRC, "tree_plan.cpp"), os.path.join(CSRC, "symbolic.cpp")]
    deps = srcs + [os.path.join(CSRC, "plan.h"), os.path.join(CSRC, "tree_plan.h"), os.path.join(CSRC, "symbolic.h")]
    if not os.path.exists(SO) or any(os.path.getmtime(f) > os.path.getmtime(SO) for f in deps):
        subprocess.run(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-o", SO] + srcs, check=True)
    L = C.CDLL(SO)
    L.plansim_factor.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                 C.c_int, C.c_int, C.c_int, C.c_void_p, C.POINTER(C.c_int), C.c_void_p, C.c_char_p, C.c_int]
    return L


def run_sim(L, A, world, relax=1, dist_NB=0, dist_min_w=0, small_max_m=0, nb=0, NB=0):
    """Interprets the launch program three times: in program order (one stream), with the look-ahead
    stream running as far ahead as its event waits allow, and with the main stream running ahead.  The
    three factors must be identical bit for bit -- a missing event dependency between the streams
    would reorder a read and a write and change the result."""
    S = sc.schol(A, relax=bool(relax))
    out = []
    for order in (0, 1, 2):
        Lx = np.empty(S.nnz)
        info = C.c_int(0)
        st = np.zeros(6, np.int64)
        err = C.create_string_buffer(2048)
        rc = L.plansim_factor(A.n, A.p.ctypes.data, A.i.ctypes.data, A.x.ctypes.data, relax, world, dist_NB, dist_min_w,
                              small_max_m, nb, NB, order, Lx.ctypes.data, C.byref(info), st.ctypes.data, err, 2048)
        assert rc == 0, f"order {order}: " + err.value.decode()
        out.append((Lx, info.value))
    for Lx, info_v in out[1:]:
        assert info_v == out[0][1] and np.array_equal(Lx, out[0][0], equal_nan=True), "stream interleavings disagree"
    Lx, info = out[0]
    return S, Lx, info, dict(zip(["n_dist", "transfers", "doubles", "ws", "events", "marks"], (int(v) for v in st)))


def check(A, Lx, tol=1e-10):
    rc, Lp, Li, ref = po.chol(A.n, A.p, A.i, A.x)
    assert rc == 0
    den = np.maximum(np.abs(ref), 1e-300)
    assert float((np.abs(Lx - ref) / den).max()) <= tol


def lap(dims):
    perm = synth.nd_perm(*dims)
    ti, tj, tx, n = synth.laplacian_triplets(*dims, perm=perm)
    return sc.triplet_to_csc_matrix(ti, tj, tx, n)


@pytest.mark.parametrize("name", ["bcsstk01", "1138_bus", "lap2d_24", "lap3d_10"])
def test_single_rank_program_matches_oracle(sim, name):
    g = np.load(os.path.join(GOLD, name + ".npz"))
    A = sc.CscMatrix(int(g["n"]), g["Ap"], g["Ai"], g["Ax"])
    # small_max_m = 24 pushes most fronts through the large-front (blocked) path as well
    for small in (0, 24):
        S, Lx, info, st = run_sim(sim, A, 1, small_max_m=small, nb=8 if small else 0, NB=16 if small else 0)
        assert info == 0 and st["transfers"] == 0
        den = np.maximum(np.abs(g["Lx"]), 1e-300)
        assert float((np.abs(Lx - g["Lx"]) / den).max()) <= 1e-10


@pytest.mark.parametrize("world", [2, 3, 4, 8])
@pytest.mark.parametrize("dims", [(40, 40, 1), (12, 12, 12), (30, 9, 7)])
def test_multi_rank_program_matches_oracle(sim, world, dims):
    A = lap(dims)
    # tiny blocks so that fronts of these small grids are really factored by rank groups
    S, Lx, info, st = run_sim(sim, A, world, dist_NB=8, dist_min_w=16, small_max_m=24, nb=8, NB=16)
    assert info == 0
    assert st["n_dist"] > 0 and st["transfers"] > 0
    check(A, Lx)


@pytest.mark.parametrize("world", [1, 2, 8])
def test_lookahead_program_and_tree_parallel_amalgamation(sim, world):
    # relax=3 keeps wide sibling separators as fronts of their own (multi-GPU mapping); the blocked
    # fronts are wide enough (w > NB) that the two-stream look-ahead program is really emitted
    A = lap((14, 14, 14))
    S, Lx, info, st = run_sim(sim, A, world, relax=3, dist_NB=8, dist_min_w=16, small_max_m=24, nb=8, NB=16)
    assert info == 0 and st["events"] > 4
    check(A, Lx)
    # wide local fronts announce their finished block columns (OP_EXPORT): the simulator checks that an announced range
    # never changes afterwards, under all three stream interleavings, and that marked fronts are announced completely
    if world == 1:
        assert st["marks"] > 4


@pytest.mark.parametrize("world", [2, 4])
def test_multi_rank_pinned_fronts_and_subtrees_only(sim, world):
    # dist_min_w huge: no front is factored by a group, split subtrees meet at fronts pinned to one rank
    A = lap((24, 24, 1))
    S, Lx, info, st = run_sim(sim, A, world, dist_NB=8, dist_min_w=10 ** 6)
    assert info == 0 and st["n_dist"] == 0 and st["transfers"] > 0
    check(A, Lx)


@pytest.mark.parametrize("world", [2, 5])
def test_multi_rank_irregular_trees(sim, world):
    for name in ("bcsstk01", "1138_bus"):
        g = np.load(os.path.join(GOLD, name + ".npz"))
        A = sc.CscMatrix(int(g["n"]), g["Ap"], g["Ai"], g["Ax"])
        S, Lx, info, st = run_sim(sim, A, world, dist_NB=4, dist_min_w=4, small_max_m=12, nb=4, NB=8)
        assert info == 0
        den = np.maximum(np.abs(g["Lx"]), 1e-300)
        assert float((np.abs(Lx - g["Lx"]) / den).max()) <= 1e-10


def test_multi_rank_default_blocking_mid_size(sim):
    # production block sizes (nb 64, NB 256, dist_NB 256 is too wide for this grid -> 64), 2 and 4 ranks
    A = lap((96, 96, 1))
    for world in (2, 4):
        S, Lx, info, st = run_sim(sim, A, world, dist_NB=64, dist_min_w=64)
        assert info == 0 and st["n_dist"] > 0
        check(A, Lx)


def test_multi_rank_not_positive_definite(sim):
    perm = synth.nd_perm(20, 20, 1)
    ti, tj, tx, n = synth.laplacian_triplets(20, 20, 1, perm)
    tx = tx.copy(); tx[n - 1] = -1.0
    A = sc.triplet_to_csc_matrix(ti, tj, tx, n)
    S, Lx, info, st = run_sim(sim, A, 2, dist_NB=8, dist_min_w=8, small_max_m=16, nb=8, NB=16)
    assert info == int(np.where(perm == n - 1)[0][0]) + 1     # 1-based column of the negative diagonal
