"""TEST INFRASTRUCTURE ONLY: ctypes access to the oracle (liboracle.so = plain-C restatement,
_ref/libref.so = the reference's own code compiled from /root/reference).  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "liboracle.so")
REF_SO = os.path.join(HERE, "_ref", "libref.so")


def build():
    subprocess.run(["make", "-C", HERE, "-s"], check=True, capture_output=True)


def _p(a):
    return None if a is None else a.ctypes.data


_VP = C.c_void_p
_orc = None
_ref = None


def oracle():
    global _orc
    if _orc is None:
        if not os.path.exists(ORACLE_SO):
            build()
        L = C.CDLL(ORACLE_SO)
        L.orc_triplet_to_csc.restype = C.c_int64
        L.orc_triplet_to_csc.argtypes = [C.c_int64, _VP, _VP, _VP, C.c_int, _VP, _VP, _VP]
        L.orc_etree.argtypes = [C.c_int, _VP, _VP, _VP]
        L.orc_post_order.argtypes = [C.c_int, _VP, _VP]
        L.orc_col_count.argtypes = [C.c_int, _VP, _VP, _VP, _VP, _VP]
        L.orc_ereach.argtypes = [C.c_int, _VP, _VP, _VP, C.c_int, _VP, _VP, _VP, _VP, C.c_int, _VP]
        L.orc_schol.restype = C.c_int64
        L.orc_schol.argtypes = [C.c_int, _VP, _VP, _VP, _VP, _VP]
        L.orc_chol.argtypes = [C.c_int, _VP, _VP, _VP, _VP, _VP, _VP, _VP]
        L.orc_compute_levels.argtypes = [C.c_int, _VP, _VP, _VP]
        L.orc_compute_supernodes.argtypes = [C.c_int, _VP, _VP, _VP, _VP]
        L.orc_atree.argtypes = [C.c_int, _VP, _VP, _VP, C.c_int, _VP, _VP]
        _orc = L
    return _orc


def have_ref():
    return os.path.exists(REF_SO)


def ref():
    global _ref
    if _ref is None:
        L = C.CDLL(REF_SO)
        L.ref_last_error.restype = C.c_char_p
        L.ref_triplet_to_csc.argtypes = [C.c_int, _VP, _VP, _VP, C.c_int, _VP, _VP, _VP]
        L.ref_load_mtx.argtypes = [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int), _VP, _VP, _VP]
        L.ref_etree.argtypes = [C.c_int, _VP, _VP, _VP]
        L.ref_post_order.argtypes = [C.c_int, _VP, _VP]
        L.ref_col_count.argtypes = [C.c_int, _VP, _VP, _VP, _VP, _VP]
        L.ref_ereach.argtypes = [C.c_int, _VP, _VP, _VP, C.c_int, _VP, _VP, _VP, _VP, C.c_int]
        L.ref_schol.argtypes = [C.c_int, _VP, _VP, _VP, _VP, _VP, C.POINTER(C.c_double)]
        L.ref_chol.argtypes = [C.c_int, _VP, _VP, _VP, _VP, _VP, _VP, C.POINTER(C.c_double)]
        L.ref_compute_levels.argtypes = [C.c_int, _VP, _VP, _VP]
        L.ref_compute_supernodes.argtypes = [C.c_int, _VP, _VP, _VP, _VP]
        L.ref_atree.argtypes = [C.c_int, _VP, _VP, _VP, _VP, C.c_int, _VP, _VP]
        L.ref_csc_to_dense_none.argtypes = [C.c_int, _VP, _VP, _VP, _VP]
        if hasattr(L, "ref_random_sparse"):
            L.ref_random_sparse.argtypes = [C.c_int, C.c_double, C.c_int, _VP, _VP, _VP]
            L.ref_random_sparse_S.argtypes = [C.c_int, _VP, _VP, _VP, C.c_int, _VP]
            L.ref_csc_from_A_S.argtypes = [C.c_int, _VP, _VP, _VP, _VP, _VP, _VP, _VP]
            L.ref_schol_etree.argtypes = [C.c_int, _VP, _VP, _VP, _VP]
        _ref = L
    return _ref


# ---- plain-C oracle ---------------------------------------------------------------------------
def triplet_to_csc(ti, tj, tx, n):
    ti = np.ascontiguousarray(ti, np.int32); tj = np.ascontiguousarray(tj, np.int32)
    tx = np.ascontiguousarray(tx, np.float64)
    nz = len(ti)
    Ap = np.empty(n + 1, np.int32); Ai = np.empty(max(nz, 1), np.int32); Ax = np.empty(max(nz, 1), np.float64)
    m = oracle().orc_triplet_to_csc(nz, _p(ti), _p(tj), _p(tx), n, _p(Ap), _p(Ai), _p(Ax))
    assert m >= 0
    return Ap, Ai[:m].copy(), Ax[:m].copy()


def etree(n, Ap, Ai):
    parent = np.empty(n, np.int32)
    oracle().orc_etree(n, _p(Ap), _p(Ai), _p(parent))
    return parent


def post_order(parent):
    post = np.empty(len(parent), np.int32)
    oracle().orc_post_order(len(parent), _p(parent), _p(post))
    return post


def col_count(n, Ap, Ai, parent, post):
    cc = np.empty(n, np.int32)
    oracle().orc_col_count(n, _p(Ap), _p(Ai), _p(parent), _p(post), _p(cc))
    return cc


def schol(n, Ap, Ai, want_Li=True):
    """-> parent, Lp (int64), Li (int32 or None)"""
    parent = np.empty(n, np.int32); Lp = np.empty(n + 1, np.int64)
    nnz = oracle().orc_schol(n, _p(Ap), _p(Ai), _p(parent), _p(Lp), None)
    assert nnz >= 0
    Li = None
    if want_Li:
        Li = np.empty(nnz, np.int32)
        oracle().orc_schol(n, _p(Ap), _p(Ai), _p(parent), _p(Lp), _p(Li))
    return parent, Lp, Li


def chol(n, Ap, Ai, Ax, parent=None, Lp=None):
    """-> rc (0 ok, k+1 = not positive definite at column k), Lp, Li, Lx"""
    if parent is None or Lp is None:
        parent, Lp, _ = schol(n, Ap, Ai, want_Li=False)
    nnz = int(Lp[n])
    Li = np.empty(nnz, np.int32); Lx = np.zeros(nnz, np.float64)
    Ax = np.ascontiguousarray(Ax, np.float64)
    rc = oracle().orc_chol(n, _p(Ap), _p(Ai), _p(Ax), _p(parent), _p(Lp), _p(Li), _p(Lx))
    return rc, Lp, Li, Lx


def compute_levels(parent):
    n = len(parent)
    lp = np.empty(n + 1, np.int32); ln = np.empty(n, np.int32)
    nl = oracle().orc_compute_levels(n, _p(parent), _p(lp), _p(ln))
    return [ln[lp[l]:lp[l + 1]].tolist() for l in range(nl)]


def compute_supernodes(parent, Lp):
    n = len(parent)
    Lp = np.ascontiguousarray(Lp, np.int64)
    sn_id = np.empty(n, np.int32); st = np.empty(n + 1, np.int32)
    ns = oracle().orc_compute_supernodes(n, _p(parent), _p(Lp), _p(sn_id), _p(st))
    return sn_id, st[:ns + 1].copy()


def atree(Lp, Li, sn_id, sn_start):
    ns = len(sn_start) - 1
    Lp = np.ascontiguousarray(Lp, np.int64)
    sp = np.empty(ns, np.int32)
    oracle().orc_atree(len(sn_id), _p(Lp), _p(Li), _p(sn_id), ns, _p(sn_start), _p(sp))
    return sp


# ---- the reference's own code -----------------------------------------------------------------
def ref_load_mtx(path):
    n = C.c_int(0); nz = C.c_int(0)
    assert ref().ref_load_mtx(path.encode(), C.byref(n), C.byref(nz), None, None, None) == 0, ref().ref_last_error()
    Ap = np.empty(n.value + 1, np.int32); Ai = np.empty(nz.value, np.int32); Ax = np.empty(nz.value, np.float64)
    ref().ref_load_mtx(path.encode(), C.byref(n), C.byref(nz), _p(Ap), _p(Ai), _p(Ax))
    return n.value, Ap, Ai, Ax


def ref_triplet_to_csc(ti, tj, tx, n):
    ti = np.ascontiguousarray(ti, np.int32); tj = np.ascontiguousarray(tj, np.int32)
    tx = np.ascontiguousarray(tx, np.float64)
    nz = len(ti)
    Ap = np.empty(n + 1, np.int32); Ai = np.empty(max(nz, 1), np.int32); Ax = np.empty(max(nz, 1), np.float64)
    m = ref().ref_triplet_to_csc(nz, _p(ti), _p(tj), _p(tx), n, _p(Ap), _p(Ai), _p(Ax))
    return Ap, Ai[:m].copy(), Ax[:m].copy()


def ref_schol(n, Ap, Ai, want_Li=True):
    parent = np.empty(n, np.int32); Lp = np.empty(n + 1, np.int32)
    sec = C.c_double(0)
    nnz = ref().ref_schol(n, _p(Ap), _p(Ai), _p(parent), _p(Lp), None, C.byref(sec))
    Li = None
    if want_Li:
        Li = np.empty(nnz, np.int32)
        ref().ref_schol(n, _p(Ap), _p(Ai), _p(parent), _p(Lp), _p(Li), C.byref(sec))
    return parent, Lp, Li, sec.value


def ref_chol(n, Ap, Ai, Ax, nnzL):
    """-> rc, Lp, Li, Lx, seconds, error string"""
    Lp = np.empty(n + 1, np.int32); Li = np.empty(nnzL, np.int32); Lx = np.empty(nnzL, np.float64)
    sec = C.c_double(0)
    Ax = np.ascontiguousarray(Ax, np.float64)
    rc = ref().ref_chol(n, _p(Ap), _p(Ai), _p(Ax), _p(Lp), _p(Li), _p(Lx), C.byref(sec))
    err = ref().ref_last_error().decode() if rc else ""
    return rc, Lp, Li, Lx, sec.value, err
