"""Host orderings (csrc/ordering.cpp; SURVEY.md 8f-2 -- additive, the reference ships no ordering):
the geometric nested dissection of the benchmark grids, approximate minimum degree, and the
symmetric permutation of an upper-triangle CSC.  All CPU; the numeric launch program of an
AMD-permuted matrix is replayed by the plan simulator against the oracle."""
import os

import numpy as np
import pytest
import scipy.sparse as sp

import sparsecholesky_b200 as sc
from oracle import pyoracle as po
from sparsecholesky_b200 import synth
from sparsecholesky_b200._lib import lib, ptr

from .test_plan_sim import run_sim, sim  # noqa: F401  (fixture)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DATA = os.path.join(ROOT, "tests", "data")


def full(A):
    M = sp.csc_matrix((A.x, A.i, A.p), shape=(A.n, A.n))
    return (M + sp.triu(M, 1).T).tocsc()


def grid(*dims):
    ti, tj, tx, n = synth.laplacian_triplets(*dims)
    return sc.triplet_to_csc_matrix(ti, tj, tx, n)


def from_dense_pattern(P):
    ti, tj = np.nonzero(np.triu(P))
    return sc.triplet_to_csc_matrix(ti.astype(np.int32), tj.astype(np.int32), np.ones(ti.size), P.shape[0])


@pytest.mark.parametrize("dims", [(1, 1, 1), (2, 1, 1), (5, 1, 1), (7, 5, 1), (4, 4, 1), (9, 9, 9), (33, 17, 1), (6, 11, 4)])
def test_nd_perm_grid_matches_numpy_statement(dims):
    a, b = synth.nd_perm(*dims), synth.nd_perm_py(*dims)
    assert np.array_equal(a, b)
    assert np.array_equal(np.sort(a), np.arange(a.size))


def test_nd_perm_check_values_of_the_survey():
    # 4x4x1 -> perm ends 2 6 10 14 (the x = 2 separator); 1024^2 starts 0 1 1024 1025 ...
    assert synth.nd_perm(4, 4, 1)[-4:].tolist() == [2, 6, 10, 14]
    assert synth.nd_perm(1024, 1024, 1)[:8].tolist() == [0, 1, 1024, 1025, 2048, 2049, 3072, 3073]
    A = sc.triplet_to_csc_matrix(*synth.laplacian_triplets(4, 4, 1, perm=synth.nd_perm(4, 4, 1)))
    S = sc.schol(A, relax=False)
    assert S.nnz == 64 and int((S.colcount().astype(np.int64) ** 2).sum()) == 282


def test_permute_upper_matches_scipy_and_map():
    rng = np.random.default_rng(5)
    n = 60
    P = rng.random((n, n)) < 0.08
    P = P | P.T | np.eye(n, dtype=bool)
    ti, tj = np.nonzero(np.triu(P))
    tx = rng.standard_normal(ti.size)
    A = sc.triplet_to_csc_matrix(ti.astype(np.int32), tj.astype(np.int32), tx, n)
    perm = rng.permutation(n).astype(np.int32)
    B, mp = sc.permute(A, perm, return_map=True)
    want = full(A)[perm][:, perm]                        # (P A P^T)[a, b] = A[perm[a], perm[b]]
    assert (abs(full(B) - want)).max() == 0.0
    assert np.all(np.diff(B.p) >= 0) and B.p[-1] == A.p[-1]
    for j in range(n):
        r = B.i[B.p[j]:B.p[j + 1]]
        assert np.all(np.diff(r) > 0) and (r.size == 0 or r[-1] <= j)
    assert np.array_equal(B.x[mp], A.x)                  # the map carries further value sets over the pattern
    # identity permutation is the identity
    I = sc.permute(A, np.arange(n, dtype=np.int32))
    assert np.array_equal(I.p, A.p) and np.array_equal(I.i, A.i) and np.array_equal(I.x, A.x)


def test_permute_rejects_a_non_permutation():
    A = grid(3, 3, 1)
    bad = np.zeros(A.n, np.int32)
    out = [np.empty(A.n + 1, np.int32), np.empty(A.i.size, np.int32), np.empty(A.i.size)]
    rc = lib().spchol_permute_upper(A.n, ptr(A.p), ptr(A.i), ptr(A.x), ptr(bad), ptr(out[0]), ptr(out[1]), ptr(out[2]), None)
    assert rc != 0 and b"permutation" in lib().spchol_last_error()


# nnz(L) under AMD must stay close to what SuperLU's multiple-minimum-degree gives on the same pattern
# (measured: 493, 3269, 63266, 864658) and far below the natural order (877, 38312, 262207, 3055619)
@pytest.mark.parametrize("name,bound,natural", [("bcsstk01", 520, 877), ("1138_bus", 3500, 38312),
                                                ("lap2d_64", 70000, 262207), ("lap3d_20", 900000, 3055619)])
def test_amd_fill(name, bound, natural):
    if name.startswith("lap2d"):
        A = grid(64, 64, 1)
    elif name.startswith("lap3d"):
        A = grid(20, 20, 20)
    else:
        A = sc.load_matrix_market_to_csc(os.path.join(DATA, name + ".mtx"))
    assert sc.schol(A).nnz == natural
    perm = sc.order_amd(A)
    assert np.array_equal(np.sort(perm), np.arange(A.n))
    B = sc.permute(A, perm)
    S = sc.schol(B)
    assert S.nnz <= bound
    # the ordering is composed with an etree postorder: every parent follows its children directly or later,
    # and a postordered tree has parent[j] > j with subtrees contiguous
    parent = S.parent
    post = sc.post_order(parent)
    assert np.array_equal(post, np.arange(A.n))


def test_amd_degenerate_patterns():
    # diagonal, one dense row (arrow), fully dense, two components, n = 1
    n = 40
    for P in (np.eye(n, dtype=bool),
              np.eye(n, dtype=bool) | (np.arange(n)[:, None] == 0) | (np.arange(n)[None, :] == 0),
              np.ones((n, n), dtype=bool),
              np.kron(np.eye(2, dtype=bool), np.ones((n // 2, n // 2), dtype=bool)),
              np.ones((1, 1), dtype=bool)):
        A = from_dense_pattern(P)
        perm = sc.order_amd(A)
        assert np.array_equal(np.sort(perm), np.arange(A.n))
        S = sc.schol(sc.permute(A, perm))
        dense = np.zeros(P.shape, bool)
        dense[np.tril_indices(P.shape[0])] = True
        assert S.nnz <= int(dense.sum())
    # arrow matrix: eliminating the hub last gives no fill at all
    arrow = from_dense_pattern(np.eye(n, dtype=bool) | (np.arange(n)[:, None] == 0) | (np.arange(n)[None, :] == 0))
    assert sc.schol(arrow).nnz == n * (n + 1) // 2
    assert sc.schol(sc.permute(arrow, sc.order_amd(arrow))).nnz == 2 * n - 1


def test_amd_dense_rows_are_ordered_last():
    # a 30 x 30 grid plus one variable coupled to everything: more than 10 sqrt(n) entries -> deferred
    g = grid(30, 30, 1)
    n = g.n + 1
    M = sp.lil_matrix((n, n))
    M[:g.n, :g.n] = sp.csc_matrix((g.x, g.i, g.p), shape=(g.n, g.n))
    M[0, 1:] = -0.001                                     # variable 0 touches every column
    M[n - 1, n - 1] = 4.0
    T = sp.triu(M.tocsc()).tocoo()
    A = sc.triplet_to_csc_matrix(T.row.astype(np.int32), T.col.astype(np.int32), T.data, n)
    perm = sc.order_amd(A)
    assert np.array_equal(np.sort(perm), np.arange(n))
    assert int(np.where(perm == 0)[0][0]) >= n - 2        # last up to the etree postorder of the final chain


@pytest.mark.parametrize("name", ["bcsstk01", "1138_bus"])
def test_amd_permuted_launch_program_matches_oracle(sim, name):  # noqa: F811
    A = sc.load_matrix_market_to_csc(os.path.join(DATA, name + ".mtx"))
    B = sc.permute(A, sc.order_amd(A))
    rc, Lp, Li, ref = po.chol(B.n, B.p, B.i, B.x)
    assert rc == 0
    S = sc.schol(B)
    assert np.array_equal(S.p, Lp) and np.array_equal(S.i, Li)           # pattern bit-exact on the permuted matrix
    for world, kw in ((1, {}), (1, dict(small_max_m=12, nb=4, NB=8)), (3, dict(dist_NB=4, dist_min_w=4, small_max_m=12, nb=4, NB=8))):
        S2, Lx, info, st = run_sim(sim, B, world, **kw)
        assert info == 0
        assert float((np.abs(Lx - ref) / np.maximum(np.abs(ref), 1e-300)).max()) <= 1e-10
