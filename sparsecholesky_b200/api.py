"""Python mirror of the reference's public host API, bound to the C-ABI (include/spchol.h).

Names, argument meaning and error behaviour follow the reference so the parity tests read like
its own tests (reference include/chol.hpp, include/mtx_reader.hpp):

    A = triplet_to_csc_matrix(ti, tj, tx, n)        # chol.hpp:309
    A = load_matrix_market_to_csc(path)             # mtx_reader.hpp:17
    S = schol(A)                                    # chol.hpp:874
    L = chol(A, S).value()   /   chol(A)            # README.md:26-27 / chol.hpp:750

The numeric phase runs only on the GPU through libspchol.so; there is no CPU fallback.
"""
import ctypes as C
import enum

import numpy as np

from ._lib import SpcholError, check, lib, ptr


class Sym(enum.Enum):           # reference `enum class sym` (chol.hpp:26)
    none = 0
    upper = 1
    lower = 2


class CscMatrix:
    """csc_matrix<double, S> (chol.hpp:134-299): p (n+1, int32), i (nnz, int32), x (nnz, f64)."""

    def __init__(self, n, p, i, x, sym=Sym.upper, m=None):
        self.n = int(n)
        self.m = int(n if m is None else m)
        self.p = np.ascontiguousarray(p, dtype=np.int32)
        self.i = np.ascontiguousarray(i, dtype=np.int32)
        self.x = np.ascontiguousarray(x, dtype=np.float64)
        self.sym = sym

    def rows(self): return self.m
    def cols(self): return self.n
    def size(self): return self.n
    def capacity(self): return int(self.i.shape[0])

    def find_index(self, i, j):                      # chol.hpp:83-95
        lo, hi = self.p[j], self.p[j + 1]
        k = lo + int(np.searchsorted(self.i[lo:hi], i))
        return k if k < hi and self.i[k] == i else -1

    def __getitem__(self, ij):                       # symmetry-aware access (chol.hpp:140-150, :230)
        i, j = ij
        if self.sym is Sym.upper and j < i: i, j = j, i
        if self.sym is Sym.lower and i < j: i, j = j, i
        k = self.find_index(i, j)
        return 0.0 if k < 0 else float(self.x[k])


class Expected:
    """std::expected<csc_matrix, std::string> (chol.hpp:750)."""

    def __init__(self, value=None, error=None):
        self._v, self._e = value, error

    def has_value(self): return self._e is None
    def __bool__(self): return self._e is None
    def error(self): return self._e

    def value(self):
        if self._e is not None:
            raise SpcholError(self._e)
        return self._v


class _Handle:
    def __init__(self, h, destroy):
        self.h, self._destroy = h, destroy

    def __del__(self):
        try:
            if self.h:
                self._destroy(self.h)
        except Exception:
            pass
        self.h = None


class SChol:
    """Reference SChol (chol.hpp:99-132): L pattern p/i + parent.  Also carries the opaque
    symbolic handle and (lazily) the device plan so chol(A, S) reuses both."""

    def __init__(self, sym_handle, n):
        self._sym = sym_handle
        self._plans = {}
        self.n = n
        L = lib()
        self.nnz = int(L.spchol_symbolic_nnzL(sym_handle.h))
        self.flops = float(L.spchol_symbolic_flops(sym_handle.h))
        self.parent = np.empty(n, np.int32)
        check(L.spchol_symbolic_parent(sym_handle.h, ptr(self.parent)))
        self._p = self._i = None

    @property
    def p(self):
        if self._p is None:
            self._p = np.empty(self.n + 1, np.int32)
            check(lib().spchol_symbolic_Lp(self._sym.h, ptr(self._p)), "SChol.p")
        return self._p

    @property
    def p64(self):
        out = np.empty(self.n + 1, np.int64)
        check(lib().spchol_symbolic_Lp64(self._sym.h, ptr(out)))
        return out

    @property
    def i(self):
        if self._i is None:
            self._i = np.empty(self.nnz, np.int32)
            check(lib().spchol_symbolic_Li(self._sym.h, ptr(self._i)), "SChol.i")
        return self._i

    def size(self): return self.n
    def capacity(self): return self.nnz

    def colcount(self):
        out = np.empty(self.n, np.int32)
        check(lib().spchol_symbolic_colcount(self._sym.h, ptr(out)))
        return out

    def supernodes(self, relaxed=False):
        """(sn_start[ns+1], sn_parent[ns]) -- fundamental (reference rule) or relaxed fronts."""
        ns = C.c_int(0)
        st = np.empty(self.n + 1, np.int32)
        pa = np.empty(self.n, np.int32)
        check(lib().spchol_symbolic_supernodes(self._sym.h, 1 if relaxed else 0, C.byref(ns), ptr(st), ptr(pa)))
        return st[: ns.value + 1].copy(), pa[: ns.value].copy()

    def panel_layout(self):
        """Supernodal export format of a single-GPU plan: dict(w, m, ld, panel_off, lbuf_entries) per front
        (see include/spchol.h, spchol_symbolic_panel_layout)."""
        ns = self.stats()["ns_relaxed"]
        w = np.empty(ns, np.int32); m = np.empty(ns, np.int32); ld = np.empty(ns, np.int32)
        off = np.empty(ns, np.int64); tot = C.c_int64(0)
        check(lib().spchol_symbolic_panel_layout(self._sym.h, ptr(w), ptr(m), ptr(ld), ptr(off), C.byref(tot)))
        return {"w": w, "m": m, "ld": ld, "panel_off": off, "lbuf_entries": int(tot.value)}

    def front_rows(self, front, m):
        rows = np.empty(int(m), np.int32)
        check(lib().spchol_symbolic_front_rows(self._sym.h, int(front), ptr(rows)))
        return rows

    def stats(self):
        st = np.zeros(8, np.int64)
        check(lib().spchol_symbolic_stats(self._sym.h, ptr(st)))
        keys = ["ns_fund", "ns_relaxed", "sum_m", "panel_entries", "padded_zeros", "max_w", "max_m", "nlevels"]
        return dict(zip(keys, (int(v) for v in st)))

    def plan(self, device=0):
        if device not in self._plans:
            h = C.c_void_p()
            check(lib().spchol_plan_create(self._sym.h, device, C.byref(h)), "spchol_plan_create")
            self._plans[device] = _Handle(h, lib().spchol_plan_destroy)
        return self._plans[device]

    def plan_dryrun(self):
        """planner statistics without a GPU (host-only)"""
        st = np.zeros(10, np.int64); ex = np.zeros(6, np.int64)
        check(lib().spchol_plan_dryrun(self._sym.h, ptr(st), ptr(ex)))
        keys = ["ns", "nlevels", "n_small", "n_large", "lbuf_entries", "ws_entries", "n_ops", "max_m", "max_w", "device_bytes"]
        d = dict(zip(keys, (int(v) for v in st)))
        d.update(dict(zip(["ops_small", "ops_extadd", "ops_potrf", "ops_trsm", "ops_gemm", "gemm_flops"], (int(v) for v in ex))))
        return d

    def plan_stats(self, device=0):
        st = np.zeros(10, np.int64)
        check(lib().spchol_plan_stats(self.plan(device).h, ptr(st)))
        keys = ["ns", "nlevels", "n_small", "n_large", "lbuf_entries", "ws_entries", "n_ops", "max_m", "max_w", "device_bytes"]
        return dict(zip(keys, (int(v) for v in st)))


def device_count():
    return int(lib().spchol_device_count())


def triplet_to_csc_matrix(ti, tj, tx, n):
    ti = np.ascontiguousarray(ti, np.int32); tj = np.ascontiguousarray(tj, np.int32)
    tx = np.ascontiguousarray(tx, np.float64)
    assert ti.shape == tj.shape == tx.shape          # chol.hpp:310
    nz = ti.shape[0]
    Ap = np.empty(n + 1, np.int32); Ai = np.empty(max(nz, 1), np.int32); Ax = np.empty(max(nz, 1), np.float64)
    m = C.c_int64(0)
    check(lib().spchol_triplet_to_csc(nz, ptr(ti), ptr(tj), ptr(tx), n, ptr(Ap), ptr(Ai), ptr(Ax), C.byref(m)),
          "triplet_to_csc_matrix")
    return CscMatrix(n, Ap, Ai[: m.value].copy(), Ax[: m.value].copy(), Sym.upper)


def triplet_to_csc_matrix_gpu(ti, tj, tx, n, return_ms=False):
    """triplet_to_csc_matrix on the GPU (spchol_triplet_to_csc_gpu: stable radix sort by (column, row) + segmented sum of
    duplicates in input order); same result as the host path / the reference (chol.hpp:309-369)."""
    ti = np.ascontiguousarray(ti, np.int32); tj = np.ascontiguousarray(tj, np.int32); tx = np.ascontiguousarray(tx, np.float64)
    nz = len(ti)
    Ap = np.empty(n + 1, np.int32); Ai = np.empty(max(nz, 1), np.int32); Ax = np.empty(max(nz, 1), np.float64)
    m = C.c_int64(0); ms = C.c_double(0)
    check(lib().spchol_triplet_to_csc_gpu(nz, ptr(ti), ptr(tj), ptr(tx), n, ptr(Ap), ptr(Ai), ptr(Ax), C.byref(m), C.byref(ms)), "triplet_to_csc_matrix_gpu")
    A = CscMatrix(n, Ap, Ai[: m.value].copy(), Ax[: m.value].copy())
    return (A, ms.value) if return_ms else A


def load_matrix_market_to_csc(filename):
    L = lib()
    n = C.c_int(0); nz = C.c_int64(0)
    rc = L.spchol_mtx_read(filename.encode(), C.byref(n), C.byref(nz), None, None, None)
    if rc < 0:
        raise RuntimeError(L.spchol_last_error().decode())      # std::runtime_error (mtx_reader.hpp:20)
    ti = np.empty(nz.value, np.int32); tj = np.empty(nz.value, np.int32); tx = np.empty(nz.value, np.float64)
    rc = L.spchol_mtx_read(filename.encode(), C.byref(n), C.byref(nz), ptr(ti), ptr(tj), ptr(tx))
    if rc < 0:
        raise RuntimeError(L.spchol_last_error().decode())
    return triplet_to_csc_matrix(ti, tj, tx, n.value)


def load_matrix_market_to_csc_gpu(filename, return_ms=False):
    """load_matrix_market_to_csc with both halves on the GPU: the text parse (spchol_mtx_read_gpu: one thread per 256-byte
    chunk, exact decimal -> double) and the COO -> CSC conversion (spchol_triplet_to_csc_gpu).  Same result, bit for bit."""
    L = lib()
    n = C.c_int(0); nz = C.c_int64(0); ms = C.c_double(0)
    rc = L.spchol_mtx_read_gpu(filename.encode(), C.byref(n), C.byref(nz), None, None, None, None)
    if rc < 0:
        raise RuntimeError(L.spchol_last_error().decode())      # std::runtime_error (mtx_reader.hpp:20)
    ti = np.empty(nz.value, np.int32); tj = np.empty(nz.value, np.int32); tx = np.empty(nz.value, np.float64)
    rc = L.spchol_mtx_read_gpu(filename.encode(), C.byref(n), C.byref(nz), ptr(ti), ptr(tj), ptr(tx), C.byref(ms))
    if rc < 0:
        raise RuntimeError(L.spchol_last_error().decode())
    A = triplet_to_csc_matrix_gpu(ti, tj, tx, n.value)
    return (A, ms.value, (ti, tj, tx)) if return_ms else A


def order_amd(A):
    """Approximate-minimum-degree ordering of the pattern of A (upper CSC): perm[new] = old.  Additive:
    the reference has no ordering."""
    perm = np.empty(A.n, np.int32)
    check(lib().spchol_order_amd(A.n, ptr(A.p), ptr(A.i), ptr(perm)), "order_amd")
    return perm


def permute(A, perm, return_map=False):
    """triu(P A P^T) with perm[new] = old; with return_map also the position in the result of every entry of A."""
    assert A.sym is Sym.upper
    perm = np.ascontiguousarray(perm, np.int32)
    Bp = np.empty(A.n + 1, np.int32); Bi = np.empty(A.i.shape[0], np.int32); Bx = np.empty(A.i.shape[0], np.float64)
    mp = np.empty(A.i.shape[0], np.int64)
    Ax = np.ascontiguousarray(A.x, np.float64)
    check(lib().spchol_permute_upper(A.n, ptr(A.p), ptr(A.i), ptr(Ax), ptr(perm), ptr(Bp), ptr(Bi), ptr(Bx), ptr(mp)), "permute")
    B = CscMatrix(A.n, Bp, Bi, Bx, Sym.upper)
    return (B, mp) if return_map else B


def build_csc_matrix_from_pattern(pattern):               # chol.hpp:416-435
    ti, tj = [], []
    for i, cols in enumerate(pattern):
        for c in cols:
            r, cc = (i, c) if i <= c else (c, i)
            ti.append(r); tj.append(cc)
    return triplet_to_csc_matrix(ti, tj, np.ones(len(ti)), len(pattern))


def etree(A):
    parent = np.empty(A.n, np.int32)
    check(lib().spchol_etree(A.n, ptr(A.p), ptr(A.i), ptr(parent)))
    return parent


def post_order(parent):
    parent = np.ascontiguousarray(parent, np.int32)
    post = np.empty(parent.shape[0], np.int32)
    check(lib().spchol_post_order(parent.shape[0], ptr(parent), ptr(post)))
    return post


def col_count(A, parent, post):
    parent = np.ascontiguousarray(parent, np.int32); post = np.ascontiguousarray(post, np.int32)
    cc = np.empty(A.n, np.int32)
    check(lib().spchol_col_count(A.n, ptr(A.p), ptr(A.i), ptr(parent), ptr(post), ptr(cc)))
    return cc


def ereach(A, k, parent, s, w, x=None, top=None):
    """In-place on s, w (int32) and x (float64) exactly like the reference overloads (chol.hpp:725/:737)."""
    parent = np.ascontiguousarray(parent, np.int32)
    top = A.n if top is None else top
    return lib().spchol_ereach(A.n, ptr(A.p), ptr(A.i), ptr(A.x) if x is not None else None, int(k),
                               ptr(parent), ptr(s), ptr(w), ptr(x) if x is not None else None, int(top))


def compute_levels(parent):
    parent = np.ascontiguousarray(parent, np.int32)
    n = parent.shape[0]
    lp = np.empty(n + 1, np.int32); ln = np.empty(n, np.int32)
    nl = lib().spchol_compute_levels(n, ptr(parent), ptr(lp), ptr(ln))
    return [ln[lp[l]: lp[l + 1]].tolist() for l in range(nl)]


def compute_supernodes(S):
    """Returns (sn_id, supernodes) like the reference (src/chol.cpp:42): supernodes = start columns + [n]."""
    n = S.n
    Lp = S.p64
    sn_id = np.empty(n, np.int32); st = np.empty(n + 1, np.int32)
    ns = lib().spchol_compute_supernodes(n, ptr(S.parent), ptr(Lp), ptr(sn_id), ptr(st))
    return sn_id, st[: ns + 1].copy()


def atree(S, sn_id, supernodes):
    ns = len(supernodes) - 1
    sn_id = np.ascontiguousarray(sn_id, np.int32); st = np.ascontiguousarray(supernodes, np.int32)
    sp = np.empty(ns, np.int32)
    check(lib().spchol_atree(S.n, ptr(S.parent), ptr(sn_id), ns, ptr(st), ptr(sp)))
    return sp


def schol(A, relax=True):
    assert A.sym is Sym.upper
    h = C.c_void_p()
    check(lib().spchol_symbolic_create_ex(A.n, ptr(A.p), ptr(A.i), int(relax), C.byref(h)), "schol")
    return SChol(_Handle(h, lib().spchol_symbolic_destroy), A.n)


def chol(A, S=None, device=0):
    """Numeric Cholesky on the GPU.  Returns Expected[CscMatrix(sym=none)]; the error string for a
    non-positive pivot is the reference's "A is not positive definite." (chol.hpp:850)."""
    if S is None:
        S = schol(A)
    if A.n != S.n or A.capacity() != int(lib().spchol_symbolic_nnzA(S._sym.h)):
        return Expected(error="chol: A does not have the pattern S was built from")
    plan = S.plan(device)
    Lx = np.empty(S.nnz, np.float64)
    info = C.c_int(0)
    Ax = np.ascontiguousarray(A.x, np.float64)
    check(lib().spchol_factor(plan.h, ptr(Ax), ptr(Lx), C.byref(info)), "chol")
    if info.value > 0:
        return Expected(error="A is not positive definite.")
    return Expected(value=CscMatrix(A.n, S.p, S.i, Lx, Sym.none))


def residual(A, S, device=0):
    """||A - L L^T||_F / ||A||_F of the factor currently held by S's device plan (the last chol(A, S)),
    estimated on the GPU with 8 random +-1 probes (spchol_residual; works without a CSC L)."""
    out = C.c_double(0)
    Ax = np.ascontiguousarray(A.x, np.float64)
    check(lib().spchol_residual(S.plan(device).h, ptr(Ax), C.byref(out)), "residual")
    return float(out.value)


def solve(S, b, device=0):
    """x = A^-1 b with the factor held by S's device plan (the last chol(A, S)): forward and backward
    substitution on the supernodal factor on the GPU (spchol_solve).  b: (n,) or (n, nrhs)."""
    b = np.asarray(b, np.float64)
    one = b.ndim == 1
    B = np.asfortranarray(b.reshape(S.n, -1))
    X = np.empty_like(B, order="F")
    check(lib().spchol_solve(S.plan(device).h, B.ctypes.data, X.ctypes.data, B.shape[1]), "solve")
    return X[:, 0].copy() if one else X


def lbuf_to_csc(S, lbuf):
    """Values of L in the CSC order of S.p / S.i, gathered on the host from a supernodal L buffer
    (spchol_factor_lbuf / spchol_lbuf_to_host of a single-GPU plan) through the documented export format."""
    lay = S.panel_layout()
    st, _ = S.supernodes(relaxed=True)
    Lp, Li = S.p64, S.i
    Lx = np.empty(int(Lp[-1]), np.float64)
    for k in range(len(lay["w"])):
        rows = S.front_rows(k, lay["m"][k])
        base, ld = int(lay["panel_off"][k]), int(lay["ld"][k])
        for c in range(int(lay["w"][k])):
            j = int(st[k]) + c
            a, b = int(Lp[j]), int(Lp[j + 1])
            pos = np.searchsorted(rows, Li[a:b])
            Lx[a:b] = lbuf[base + c * ld + pos]
    return Lx


def chol_batched(S, Ax, device=0):
    """Config 5: numeric factorization of many matrices sharing the pattern behind S.
    Ax: (nbatch, nnz(A)) values in the CSC order of the pattern A.  Returns (Lx (nbatch, nnz(L)) in
    the CSC order of S.p / S.i, info (nbatch,) with 0 or the 1-based failing column per matrix) --
    what a loop of chol(A_b, S) over the batch would give (reference include/chol.hpp:750)."""
    Ax = np.ascontiguousarray(Ax, np.float64)
    assert Ax.ndim == 2
    nb = Ax.shape[0]
    plan = S.plan(device)
    Lx = np.empty((nb, S.nnz), np.float64)
    info = np.zeros(nb, np.int32)
    check(lib().spchol_factor_batched(plan.h, nb, ptr(Ax), ptr(Lx), ptr(info)), "chol_batched")
    return Lx, info


def csc_to_dense(M):                                       # chol.hpp:1449-1479 (column-major in the reference)
    D = np.zeros((M.m, M.n))
    for j in range(M.n):
        for k in range(M.p[j], M.p[j + 1]):
            i = M.i[k]
            D[i, j] = M.x[k]
            if M.sym is not Sym.none and i != j:
                D[j, i] = M.x[k]
    return D
