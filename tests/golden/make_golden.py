"""Generate tests/golden/*.npz by running the REFERENCE'S OWN code (oracle/_ref/libref.so, built
from /root/reference by oracle/Makefile).  Run in the build container only:
    python tests/golden/make_golden.py
The fixtures are what pins the plain-C oracle and the product's symbolic phase."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pyoracle as po  # noqa: E402
from sparsecholesky_b200 import synth  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def pattern_triplets(pattern):
    ti, tj = [], []
    for i, cols in enumerate(pattern):
        for c in cols:
            r, cc = (i, c) if i <= c else (c, i)
            ti.append(r); tj.append(cc)
    return np.array(ti, np.int32), np.array(tj, np.int32), np.ones(len(ti)), len(pattern)


def golden_for(name, n, Ap, Ai, Ax, numeric=True):
    R = po.ref()
    parent, Lp, Li, _ = po.ref_schol(n, Ap, Ai)
    post = np.empty(n, np.int32); R.ref_post_order(n, parent.ctypes.data, post.ctypes.data)
    et = np.empty(n, np.int32); R.ref_etree(n, Ap.ctypes.data, Ai.ctypes.data, et.ctypes.data)
    assert (et == parent).all()
    cc = np.empty(n, np.int32)
    R.ref_col_count(n, Ap.ctypes.data, Ai.ctypes.data, parent.ctypes.data, post.ctypes.data, cc.ctypes.data)
    sn_id = np.empty(n, np.int32); st = np.empty(n + 1, np.int32)
    ns = R.ref_compute_supernodes(n, parent.ctypes.data, Lp.ctypes.data, sn_id.ctypes.data, st.ctypes.data)
    st = st[:ns + 1].copy()
    sp = np.empty(ns, np.int32)
    R.ref_atree(n, parent.ctypes.data, Lp.ctypes.data, Li.ctypes.data, sn_id.ctypes.data, ns, st.ctypes.data, sp.ctypes.data)
    lp = np.empty(n + 1, np.int32); ln = np.empty(n, np.int32)
    nl = R.ref_compute_levels(n, parent.ctypes.data, lp.ctypes.data, ln.ctypes.data)
    alp = np.empty(ns + 1, np.int32); aln = np.empty(ns, np.int32)
    anl = R.ref_compute_levels(ns, sp.ctypes.data, alp.ctypes.data, aln.ctypes.data)
    d = dict(n=n, Ap=Ap, Ai=Ai, Ax=Ax, parent=parent, post=post, colcount=cc, Lp=Lp, Li=Li,
             sn_id=sn_id, sn_start=st, atree=sp, level_ptr=lp[:nl + 1], level_nodes=ln,
             alevel_ptr=alp[:anl + 1], alevel_nodes=aln)
    if numeric:
        rc, Lp2, Li2, Lx, _, err = po.ref_chol(n, Ap, Ai, Ax, int(Lp[n]))
        d["chol_rc"] = rc
        d["chol_err"] = np.array(err)
        if rc == 0:
            assert (Lp2 == Lp).all() and (Li2 == Li).all()
            d["Lx"] = Lx
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **d)
    print(name, "n", n, "nnzA", len(Ai), "nnzL", int(Lp[n]), "ns", ns, "levels", nl, anl)


def main():
    assert po.have_ref(), "build oracle/_ref first (make -C oracle)"
    # thesis p.26 7x7 (reference tests/test_chol.cpp:7-15)
    p7 = [[0], [1], [0, 2], [3], [0, 2, 4], [0, 1, 3, 5], [0, 2, 5, 6]]
    ti, tj, tx, n = pattern_triplets(p7)
    golden_for("thesis7", n, *po.ref_triplet_to_csc(ti, tj, tx, n), numeric=False)
    # thesis example 9 (reference src/main.cpp:272-282, commented demo)
    p9 = [[0], [0, 1], [2], [2, 3], [4], [4, 5], [3, 4, 6], [0, 1, 2, 3, 5, 7], [0, 1, 5, 6, 8]]
    ti, tj, tx, n = pattern_triplets(p9)
    golden_for("thesis9", n, *po.ref_triplet_to_csc(ti, tj, tx, n), numeric=False)
    # 3x3 of tests/test_chol.cpp:59-97 (explicit zero at (1,2) stays structural)
    ti = [0, 0, 0, 1, 1, 2]; tj = [0, 1, 2, 1, 2, 2]; tx = [4.0, 1.0, 1.0, 3.0, 0.0, 2.0]
    golden_for("test3", 3, *po.ref_triplet_to_csc(ti, tj, tx, 3))
    # not positive definite 2x2
    golden_for("notpd2", 2, *po.ref_triplet_to_csc([0, 0, 1], [0, 1, 1], [1.0, 2.0, 1.0], 2))
    # shipped matrices
    for nm in ("bcsstk01", "1138_bus"):
        n, Ap, Ai, Ax = po.ref_load_mtx(os.path.join(ROOT, "tests", "data", nm + ".mtx"))
        golden_for(nm, n, Ap, Ai, Ax)
    # small Laplacians under the ND permutation
    for nm, dims in (("lap2d_24", (24, 24, 1)), ("lap3d_10", (10, 10, 10)), ("lap2d_7x5", (7, 5, 1))):
        perm = synth.nd_perm(*dims)
        ti, tj, tx, n = synth.laplacian_triplets(*dims, perm=perm)
        golden_for(nm, n, *po.ref_triplet_to_csc(ti, tj, tx, n))


def boundary_golden():
    """outputs of the reference's remaining public helpers (random_sparse x2, csc_matrix(A, S), schol_etree) for the
    drop-in header test (tests/test_dropin_cpp.py)"""
    import ctypes as C
    R = po.ref()
    out = {}
    for n, dens in ((10, 0.25), (37, 0.1)):
        nz = R.ref_random_sparse(n, dens, 1, None, None, None)
        Ap = np.empty(n + 1, np.int32); Ai = np.empty(nz, np.int32); Ax = np.empty(nz)
        R.ref_random_sparse(n, dens, 1, Ap.ctypes.data, Ai.ctypes.data, Ax.ctypes.data)
        out[f"rs{n}_p"] = Ap; out[f"rs{n}_i"] = Ai; out[f"rs{n}_x"] = Ax
    n, Ap, Ai, Ax = po.ref_load_mtx(os.path.join(ROOT, "tests", "data", "bcsstk01.mtx"))
    parent, Lp, Li, _ = po.ref_schol(n, Ap, Ai)
    x = np.empty(int(Lp[n]))
    R.ref_random_sparse_S(n, parent.ctypes.data, Lp.ctypes.data, Li.ctypes.data, 1, x.ctypes.data)
    out["rsS_x"] = x.copy()
    R.ref_csc_from_A_S(n, Ap.ctypes.data, Ai.ctypes.data, Ax.ctypes.data, parent.ctypes.data, Lp.ctypes.data, Li.ctypes.data, x.ctypes.data)
    out["ctor_x"] = x.copy()
    Ep = np.empty(n + 1, np.int32)
    nz = R.ref_schol_etree(n, Ap.ctypes.data, Ai.ctypes.data, Ep.ctypes.data, None)
    Ei = np.empty(nz, np.int32)
    R.ref_schol_etree(n, Ap.ctypes.data, Ai.ctypes.data, Ep.ctypes.data, Ei.ctypes.data)
    out["se_p"] = Ep; out["se_i"] = Ei
    np.savez_compressed(os.path.join(OUT, "boundary.npz"), **out)
    print("boundary: random_sparse(10) nnz", len(out["rs10_i"]), "schol_etree(bcsstk01) nnz", nz)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "boundary":
        boundary_golden()
    else:
        main()
        boundary_golden()
