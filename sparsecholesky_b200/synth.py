"""Deterministic synthetic inputs of BASELINE.json's configs: 5-/7-point grid
Laplacians under the fixed geometric nested-dissection permutation, and the batched D*A*D family.
Harness code: the reference ships no ordering and no generators for these."""
import numpy as np


def nd_perm(nx, ny=1, nz=1):
    """perm[new] = old grid id (id = x + nx*(y + ny*z)); geometric nested dissection, leaf volume <= 8
    (spchol_nd_perm_grid; nd_perm_py below is the same in numpy and checks it in the tests)."""
    from ._lib import check, lib, ptr
    out = np.empty(nx * ny * nz, np.int64)
    check(lib().spchol_nd_perm_grid(nx, ny, nz, ptr(out)), "nd_perm")
    return out


def nd_perm_py(nx, ny=1, nz=1):
    out = np.empty(nx * ny * nz, np.int64)
    pos = 0

    def emit(x0, x1, y0, y1, z0, z1):
        nonlocal pos
        zz, yy, xx = np.meshgrid(np.arange(z0, z1), np.arange(y0, y1), np.arange(x0, x1), indexing="ij")
        ids = (xx + nx * (yy + ny * zz)).ravel()
        out[pos:pos + ids.size] = ids
        pos += ids.size

    stack = [("box", 0, nx, 0, ny, 0, nz)]
    while stack:
        kind, x0, x1, y0, y1, z0, z1 = stack.pop()
        dx, dy, dz = x1 - x0, y1 - y0, z1 - z0
        if dx <= 0 or dy <= 0 or dz <= 0:
            continue
        if kind == "emit" or dx * dy * dz <= 8 or max(dx, dy, dz) < 3:
            emit(x0, x1, y0, y1, z0, z1)
            continue
        # children pushed in reverse so they pop as: left, right, separator
        if dx >= dy and dx >= dz:
            m = x0 + dx // 2
            stack.append(("emit", m, m + 1, y0, y1, z0, z1))
            stack.append(("box", m + 1, x1, y0, y1, z0, z1))
            stack.append(("box", x0, m, y0, y1, z0, z1))
        elif dy >= dz:
            m = y0 + dy // 2
            stack.append(("emit", x0, x1, m, m + 1, z0, z1))
            stack.append(("box", x0, x1, m + 1, y1, z0, z1))
            stack.append(("box", x0, x1, y0, m, z0, z1))
        else:
            m = z0 + dz // 2
            stack.append(("emit", x0, x1, y0, y1, m, m + 1))
            stack.append(("box", x0, x1, y0, y1, m + 1, z1))
            stack.append(("box", x0, x1, y0, y1, z0, m))
    assert pos == out.size
    return out


def laplacian_triplets(nx, ny=1, nz=1, perm=None):
    """Upper-triangle triplets (ti <= tj not required) of the 5-/7-point Laplacian (Dirichlet),
    diag = 2*dim, off-diag = -1, relabelled by `perm` (perm[new] = old)."""
    n = nx * ny * nz
    dim = (nx > 1) + (ny > 1) + (nz > 1)
    ids = np.arange(n, dtype=np.int64)
    x = ids % nx; y = (ids // nx) % ny; z = ids // (nx * ny)
    rows = [ids]; cols = [ids]; vals = [np.full(n, 2.0 * dim)]
    for mask, step in ((x + 1 < nx, 1), (y + 1 < ny, nx), (z + 1 < nz, nx * ny)):
        a = ids[mask]
        rows.append(a); cols.append(a + step); vals.append(np.full(a.size, -1.0))
    ti = np.concatenate(rows); tj = np.concatenate(cols); tx = np.concatenate(vals)
    if perm is not None:
        inv = np.empty(n, np.int64)
        inv[perm] = np.arange(n, dtype=np.int64)
        ti = inv[ti]; tj = inv[tj]
    return ti.astype(np.int32), tj.astype(np.int32), tx, n


def splitmix64(x):
    x = np.asarray(x, dtype=np.uint64)
    with np.errstate(over="ignore"):
        x = x + np.uint64(0x9E3779B97F4A7C15)
        z = x
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return z


def batch_scaling(b, n, seed=0x4D2):
    """d_i = 1 + 0.5*u_{b,i}, u from a stateless splitmix64 counter."""
    ctr = (np.uint64(seed) << np.uint64(32)) ^ (np.uint64(b) * np.uint64(n) + np.arange(n, dtype=np.uint64))
    u = (splitmix64(ctr) >> np.uint64(11)).astype(np.float64) * (2.0 ** -53)
    return 1.0 + 0.5 * u


def batch_values(Ap, Ai, Ax, b):
    """values of A_b = D_b A D_b on the pattern of A, evaluated as (d_i * a_ij) * d_j"""
    n = len(Ap) - 1
    d = batch_scaling(b, n)
    col = np.repeat(np.arange(n), np.diff(Ap))
    return (d[Ai] * Ax) * d[col]


def graded_laplacian_triplets(nx, ny=1, nz=1, perm=None, decades=10.0, shift=1e-9, seed=7):
    """Ill-conditioned SPD test matrix with the pattern of laplacian_triplets: a weighted graph Laplacian whose edge
    weights span `decades` orders of magnitude (10^U(-decades/2, decades/2), seeded), free boundary, plus
    shift * (largest diagonal) on the diagonal so that it is (barely) positive definite.  cond(A) ~ 1e10 and more with
    the defaults -- the case the explicit-inverse panel solve of the wide fronts has to survive."""
    n = nx * ny * nz
    ids = np.arange(n, dtype=np.int64)
    x = ids % nx; y = (ids // nx) % ny; z = ids // (nx * ny)
    rng = np.random.default_rng(seed)
    rows, cols, vals = [], [], []
    diag = np.zeros(n)
    for mask, step in ((x + 1 < nx, 1), (y + 1 < ny, nx), (z + 1 < nz, nx * ny)):
        a = ids[mask]
        if a.size == 0:
            continue
        w = 10.0 ** rng.uniform(-decades / 2, decades / 2, a.size)
        rows.append(a); cols.append(a + step); vals.append(-w)
        np.add.at(diag, a, w); np.add.at(diag, a + step, w)
    diag += shift * diag.max()
    ti = np.concatenate([ids] + rows); tj = np.concatenate([ids] + cols); tx = np.concatenate([diag] + vals)
    if perm is not None:
        inv = np.empty(n, np.int64)
        inv[perm] = np.arange(n, dtype=np.int64)
        ti = inv[ti]; tj = inv[tj]
    return ti.astype(np.int32), tj.astype(np.int32), tx, n
