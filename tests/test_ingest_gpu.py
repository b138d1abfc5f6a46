"""Device-side ingest: COO -> upper CSC on the GPU (csrc/ingest.cuh: stable radix sort by (column, row),
segmented sum of duplicates) against the checker's triplet_to_csc -- the plain-C restatement of the reference's
triplet_to_csc_matrix (include/chol.hpp:309-369), and the reference itself where oracle/_ref is present.  Bit for bit."""
import numpy as np
import pytest

import sparsecholesky_b200 as sc
from oracle import pyoracle as po
from sparsecholesky_b200 import synth

pytestmark = pytest.mark.gpu


def same(A, Ap, Ai, Ax):
    assert np.array_equal(A.p, Ap) and np.array_equal(A.i, Ai)
    assert np.array_equal(A.x, Ax)                       # same additions in the same order


@pytest.mark.parametrize("n,nz,seed", [(1, 1, 0), (7, 30, 1), (100, 5000, 2), (1000, 3000, 3), (5000, 200000, 4), (70000, 400000, 5)])
def test_random_triplets_with_duplicates(n, nz, seed):
    rng = np.random.default_rng(seed)
    ti = rng.integers(0, n, nz).astype(np.int32); tj = rng.integers(0, n, nz).astype(np.int32)
    tx = rng.standard_normal(nz)
    tx[rng.integers(0, nz, max(1, nz // 10))] = 0.0      # explicit zeros stay structural (reference tests/test_chol.cpp:79)
    A = sc.triplet_to_csc_matrix_gpu(ti, tj, tx, n)
    same(A, *po.triplet_to_csc(ti, tj, tx, n))
    H = sc.triplet_to_csc_matrix(ti, tj, tx, n)           # the host path of the product
    same(A, H.p, H.i, H.x)
    if po.have_ref() and nz <= 5000:
        # the reference's std::sort leaves the order of duplicates unspecified: compare where sums of two are exact anyway
        Rp, Ri, Rx = po.ref_triplet_to_csc(ti, tj, tx, n)
        assert np.array_equal(A.p, Rp) and np.array_equal(A.i, Ri) and np.allclose(A.x, Rx, rtol=1e-14, atol=1e-300)


def test_many_duplicates_are_added_in_input_order():
    # 200 triplets on 3 slots, values of very different size: the sum depends on the order, which must be the input order
    rng = np.random.default_rng(9)
    ti = rng.integers(0, 2, 200).astype(np.int32); tj = rng.integers(0, 2, 200).astype(np.int32)
    tx = rng.standard_normal(200) * 10.0 ** rng.integers(-8, 8, 200)
    A = sc.triplet_to_csc_matrix_gpu(ti, tj, tx, 2)
    same(A, *po.triplet_to_csc(ti, tj, tx, 2))


@pytest.mark.parametrize("dims", [(64, 64, 1), (20, 20, 20), (1024, 1024, 1)])
def test_laplacians_and_speed(dims):
    perm = synth.nd_perm(*dims)
    ti, tj, tx, n = synth.laplacian_triplets(*dims, perm=perm)
    A, ms = sc.triplet_to_csc_matrix_gpu(ti, tj, tx, n, return_ms=True)
    A, ms = sc.triplet_to_csc_matrix_gpu(ti, tj, tx, n, return_ms=True)     # second call: work space allocated, kernels loaded
    H = sc.triplet_to_csc_matrix(ti, tj, tx, n)
    same(A, H.p, H.i, H.x)
    print(f"device ingest {dims}: {len(ti)} triplets in {ms:.2f} ms")
    if dims == (1024, 1024, 1):
        assert ms <= 20.0                                 # round-1 verdict target for config 3


def test_bad_index_is_refused():
    with pytest.raises(Exception):
        sc.triplet_to_csc_matrix_gpu([0, 5], [0, 1], [1.0, 2.0], 3)
