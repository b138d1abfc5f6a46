"""Host logic of the product (C-ABI, no GPU): ingest + symbolic phase must reproduce the
reference's S.p / S.i / S.parent / supernodes / atree / levels bit-exactly (golden vectors made by
the reference's own code), and the library must export every symbol include/spchol.h declares."""
import ctypes
import glob
import os
import re

import numpy as np
import pytest

import sparsecholesky_b200 as sc
from sparsecholesky_b200 import _lib, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
NAMES = sorted(os.path.basename(f)[:-4] for f in glob.glob(os.path.join(GOLD, "*.npz")) if not os.path.basename(f).startswith(("large_", "boundary")))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "spchol.h")).read()
    declared = set(re.findall(r"\b(spchol_[A-Za-z0-9_]+)\s*\(", hdr))
    assert len(declared) > 30
    L = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(L, name), name
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)


@pytest.mark.parametrize("name", NAMES)
def test_symbolic_matches_reference_golden(name):
    g = np.load(os.path.join(GOLD, name + ".npz"))
    n = int(g["n"])
    A = sc.CscMatrix(n, g["Ap"], g["Ai"], g["Ax"])
    parent = sc.etree(A)
    assert (parent == g["parent"]).all()
    post = sc.post_order(parent)
    assert (post == g["post"]).all()
    assert (sc.col_count(A, parent, post) == g["colcount"]).all()
    for relax in (False, True):
        S = sc.schol(A, relax=relax)
        assert (S.parent == g["parent"]).all()
        assert (S.p == g["Lp"]).all() and (S.i == g["Li"]).all()
        assert S.flops == float((g["colcount"].astype(np.float64) ** 2).sum())
        sn_id, st = sc.compute_supernodes(S)
        assert (sn_id == g["sn_id"]).all() and (st == g["sn_start"]).all()
        assert (sc.atree(S, sn_id, st) == g["atree"]).all()
        fst, fpa = S.supernodes(relaxed=False)
        assert (fst == g["sn_start"]).all() and (fpa == g["atree"]).all()
    lv = sc.compute_levels(parent)
    glv = [g["level_nodes"][g["level_ptr"][l]:g["level_ptr"][l + 1]].tolist() for l in range(len(g["level_ptr"]) - 1)]
    assert lv == glv
    alv = sc.compute_levels(g["atree"])
    galv = [g["alevel_nodes"][g["alevel_ptr"][l]:g["alevel_ptr"][l + 1]].tolist() for l in range(len(g["alevel_ptr"]) - 1)]
    assert alv == galv


def test_reference_unit_tests_restated():
    # tests/test_chol.cpp:6-25 EliminationTree
    pattern = [[0], [1], [0, 2], [3], [0, 2, 4], [0, 1, 3, 5], [0, 2, 5, 6]]
    A = sc.build_csc_matrix_from_pattern(pattern)
    parent = sc.etree(A)
    assert parent.tolist() == [2, 5, 4, 5, 5, 6, -1]
    # tests/test_chol.cpp:27-57 ColumnReach (both overloads)
    n = A.size()
    w = np.full(n, -1, np.int32); s = np.zeros(n, np.int32); x = np.zeros(n)
    sc.ereach(A, 5, parent, s, w, x, n)
    assert s.tolist() == [3, 1, 0, 2, 4, 5, 6]
    w[:] = -1                      # the reference re-runs on the already-marked w; same s stays
    sc.ereach(A, 5, parent, s, w, None, n)
    assert s.tolist() == [3, 1, 0, 2, 4, 5, 6]


def test_triplet_rules():
    # mirror to upper, sum duplicates, keep explicit zeros (chol.hpp:321-346; tests/test_chol.cpp:79)
    A = sc.triplet_to_csc_matrix([0, 1, 2, 2, 1, 0, 2], [0, 1, 2, 0, 0, 1, 1], [4.0, 3.0, 2.0, 1.0, 0.5, 0.5, 0.0], 3)
    assert A.p.tolist() == [0, 1, 3, 6]
    assert A.i.tolist() == [0, 0, 1, 0, 1, 2]
    assert A.x.tolist() == [4.0, 1.0, 3.0, 1.0, 0.0, 2.0]
    assert A[1, 0] == 1.0 and A[0, 1] == 1.0 and A[2, 1] == 0.0
    with pytest.raises(sc.api.SpcholError):
        sc.triplet_to_csc_matrix([0, 5], [0, 0], [1.0, 1.0], 3)


def test_mtx_reader():
    A = sc.load_matrix_market_to_csc(os.path.join(ROOT, "tests", "data", "bcsstk01.mtx"))
    g = np.load(os.path.join(GOLD, "bcsstk01.npz"))
    assert A.n == 48 and (A.p == g["Ap"]).all() and (A.i == g["Ai"]).all() and np.array_equal(A.x, g["Ax"])
    with pytest.raises(RuntimeError, match="Could not open file"):
        sc.load_matrix_market_to_csc("/nonexistent/file.mtx")


def test_mtx_reader_token_grammar(tmp_path):
    # same grammar as the reference's `ifstream >>` loop (mtx_reader.hpp:17-62): leading % lines, then
    # whitespace-separated tokens; signs, exponents, CRLF, blank lines, several entries per line
    text = ("%%MatrixMarket matrix coordinate real symmetric\r\n% a comment\r\n%\r\n"
            "  4 4   6\r\n1 1 +2.5\r\n\r\n2 1 -1e-1   2 2 3.\n3 3 1E+2\n4 3 .5\n4 4 1e400\n")
    f = tmp_path / "a.mtx"
    f.write_bytes(text.encode())
    A = sc.load_matrix_market_to_csc(str(f))
    assert A.n == 4 and A.p.tolist() == [0, 1, 3, 4, 6] and A.i.tolist() == [0, 0, 1, 2, 2, 3]
    assert A.x[:5].tolist() == [2.5, -0.1, 3.0, 100.0, 0.5] and np.isinf(A.x[5])
    # values parse to the same doubles as the C library's strtod (what operator>> uses)
    vals = ["0.1", "1138.000123456789", "-4.9406564584124654e-324", "1.7976931348623157e308", "123456789012345678", "6.02214076e23"]
    g = tmp_path / "b.mtx"
    g.write_text("6 6 6\n" + "".join(f"{k + 1} {k + 1} {v}\n" for k, v in enumerate(vals)))
    B = sc.load_matrix_market_to_csc(str(g))
    assert B.x.tolist() == [float(v) for v in vals]
    # a long comment block (header beyond the first read of the sizes-only call)
    h = tmp_path / "c.mtx"
    h.write_text("% x\n" * 40000 + "2 2 2\n1 1 1\n2 2 1\n")
    assert sc.load_matrix_market_to_csc(str(h)).n == 2
    for bad, msg in (("3 3 2\n1 1 1.0\n", "truncated"), ("3 4 1\n1 1 1.0\n", "square"), ("% only\n", "size line"),
                     ("2 2 1\n1 x 1.0\n", "truncated")):
        b = tmp_path / "bad.mtx"
        b.write_text(bad)
        with pytest.raises(RuntimeError, match=msg):
            sc.load_matrix_market_to_csc(str(b))


def test_survey_check_values_large():
    # SURVEY.md §8(d) check values for the generator + symbolic at a size the reference needs minutes for
    perm = synth.nd_perm(4, 4, 1)
    assert perm[-4:].tolist() == [2, 6, 10, 14]
    ti, tj, tx, n = synth.laplacian_triplets(4, 4, 1, perm)
    S = sc.schol(sc.triplet_to_csc_matrix(ti, tj, tx, n), relax=False)
    assert S.nnz == 64 and S.flops == 282 and S.stats()["ns_fund"] == 11
    perm = synth.nd_perm(256, 256, 1)
    ti, tj, tx, n = synth.laplacian_triplets(256, 256, 1, perm)
    S = sc.schol(sc.triplet_to_csc_matrix(ti, tj, tx, n))
    assert S.nnz == 2332245                      # BASELINE.md §2, 2D 256^2 ND
    perm = synth.nd_perm(32, 32, 32)
    ti, tj, tx, n = synth.laplacian_triplets(32, 32, 32, perm)
    S = sc.schol(sc.triplet_to_csc_matrix(ti, tj, tx, n))
    assert S.nnz == 8130857 and S.flops == 5708347727.0     # BASELINE.md §2, 3D 32^3 ND
    st = S.stats()
    # relaxed fronts: fewer, never lose a column, padded storage accounted
    assert st["ns_relaxed"] < st["ns_fund"] and st["panel_entries"] >= S.nnz


@pytest.mark.parametrize("relax", [0, 1, 3])
def test_supernodal_export_format(relax):
    """The documented layout of the supernodal L buffer (spchol_symbolic_panel_layout / _front_rows): panels tile
    the buffer exactly as the planner lays it out, and every entry of the reference's pattern has a slot."""
    g = np.load(os.path.join(GOLD, "1138_bus.npz"))
    A = sc.CscMatrix(int(g["n"]), g["Ap"], g["Ai"], g["Ax"])
    S = sc.schol(A, relax=relax)
    lay = S.panel_layout()
    st, pa = S.supernodes(relaxed=True)
    ns = len(pa)
    assert lay["lbuf_entries"] == S.plan_dryrun()["lbuf_entries"]
    assert (lay["w"] == np.diff(st)).all() and (lay["ld"] >= lay["m"]).all() and (lay["ld"] % 2 == 0).all()
    off = np.concatenate([[0], np.cumsum(lay["ld"].astype(np.int64) * lay["w"])])
    assert (lay["panel_off"] == off[:-1]).all() and off[-1] == lay["lbuf_entries"]
    Lp, Li = S.p, S.i
    for k in range(ns):
        rows = S.front_rows(k, lay["m"][k])
        assert (np.diff(rows) > 0).all() and (rows[: lay["w"][k]] == np.arange(st[k], st[k + 1])).all()
        for j in range(st[k], st[k + 1]):
            col = Li[Lp[j]:Lp[j + 1]]
            pos = np.searchsorted(rows, col)
            assert (pos < len(rows)).all() and (rows[pos] == col).all()      # every true entry has a slot
        if pa[k] >= 0:                                                        # below-rows live in the parent front
            prow = S.front_rows(pa[k], lay["m"][pa[k]])
            assert np.isin(rows[lay["w"][k]:], prow).all()


def test_numeric_path_fails_loudly_without_gpu():
    if sc.device_count() > 0:
        pytest.skip("GPU present")
    A = sc.triplet_to_csc_matrix([0, 0, 1], [0, 1, 1], [4.0, 1.0, 3.0], 2)
    with pytest.raises(sc.api.SpcholError, match="no CUDA device"):
        sc.chol(A)


@pytest.mark.parametrize("name", ["bcsstk01", "1138_bus", "lap3d_10"])
def test_gather_csc64_from_panels(name):
    """spchol_symbolic_gather_csc64 (the 64-bit CSC behind chol64 in the C++ header): pattern == the reference's S.p / S.i,
    values gathered from a supernodal panel buffer filled with the reference's L through the documented export layout."""
    import ctypes as C
    from sparsecholesky_b200._lib import check, lib, ptr
    g = np.load(os.path.join(GOLD, name + ".npz"))
    A = sc.CscMatrix(int(g["n"]), g["Ap"], g["Ai"], g["Ax"])
    S = sc.schol(A)
    n, nnz = A.n, S.nnz
    lay = S.panel_layout()
    st, _ = S.supernodes(relaxed=True)
    lbuf = np.full(lay["lbuf_entries"], np.nan)
    Lp, Li, Lx = g["Lp"], g["Li"], g["Lx"]
    for k in range(len(lay["w"])):                       # scatter the reference's L into the panels (padding stays NaN)
        rows = S.front_rows(k, lay["m"][k])
        for c in range(int(lay["w"][k])):
            j = int(st[k]) + c
            a, b = int(Lp[j]), int(Lp[j + 1])
            pos = np.searchsorted(rows, Li[a:b])
            lbuf[int(lay["panel_off"][k]) + c * int(lay["ld"][k]) + pos] = Lx[a:b]
    p64 = np.empty(n + 1, np.int64); i32 = np.empty(nnz, np.int32); x = np.empty(nnz)
    check(lib().spchol_symbolic_gather_csc64(S._sym.h, ptr(lbuf), ptr(p64), ptr(i32), ptr(x)), "gather_csc64")
    assert np.array_equal(p64, Lp.astype(np.int64)) and np.array_equal(i32, Li)
    assert np.array_equal(x, Lx)                           # only true entries were read: no NaN came through
    check(lib().spchol_symbolic_gather_csc64(S._sym.h, None, ptr(p64), ptr(i32), None), "gather_csc64 pattern only")
    assert int(lib().spchol_symbolic_nnzA(S._sym.h)) == A.capacity()
