"""Full-size value-parity fixtures: ONE run of the REFERENCE'S OWN chol() (oracle/_ref/libref.so, built from
/root/reference by oracle/Makefile) on BASELINE.json's config 3 (2D 1024^2) and on the largest 3D cube the
reference finishes in minutes (64^3), under the fixed geometric nested-dissection permutation (SURVEY.md
§8d).  The reference needs ~19 min (1024^2) / ~10 min (64^3) of one CPU core, so the result is reduced to
a small fixture here, in the build container:

    python tests/golden/make_golden_large.py lap2d_1024      # or lap3d_64, or both

Fixture = sha256 of the reference's L.p / L.i (pattern, bit-exact), K sampled L.x entries at positions
drawn from a seeded generator (np.random.default_rng(seed).integers(0, nnzL, K)), and per-chunk sums
(sum x, sum |x|, sum x^2 over 4096 equal chunks of L.x) so every entry of L is covered by a checksum.
tests/test_chol_gpu.py::test_fullsize_vs_reference compares the GPU factor with it."""
import hashlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pyoracle as po  # noqa: E402
from sparsecholesky_b200 import synth  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
CASES = {"lap2d_1024": (1024, 1024, 1), "lap3d_64": (64, 64, 64), "lap3d_32": (32, 32, 32), "lap2d_256": (256, 256, 1)}
K = 1 << 17
SEED = 20261105
NCHUNK = 4096


def chunk_stats(x, nchunk=NCHUNK):
    """(sum x, sum |x|, sum x^2) over nchunk equal slices of x (np.array_split boundaries)"""
    n = x.size
    bounds = (np.arange(nchunk + 1, dtype=np.int64) * n) // nchunk
    s = np.add.reduceat(x, bounds[:-1])
    a = np.add.reduceat(np.abs(x), bounds[:-1])
    q = np.add.reduceat(x * x, bounds[:-1])
    return bounds, s, a, q


def make(name):
    dims = CASES[name]
    assert po.have_ref(), "build oracle/_ref first (make -C oracle)"
    perm = synth.nd_perm_py(*dims)
    ti, tj, tx, n = synth.laplacian_triplets(*dims, perm=perm)
    Ap, Ai, Ax = po.ref_triplet_to_csc(ti, tj, tx, n)
    # nnz(L) from the plain-C oracle's near-linear column counts (the reference's schol is O(n^2))
    parent = po.etree(n, Ap, Ai)
    cc = po.col_count(n, Ap, Ai, parent, po.post_order(parent))
    nnzL = int(cc.astype(np.int64).sum())
    assert nnzL < 2 ** 31
    t0 = time.time()
    rc, Lp, Li, Lx, sec, err = po.ref_chol(n, Ap, Ai, Ax, nnzL)
    assert rc == 0, err
    assert int(Lp[n]) == nnzL
    pos = np.random.default_rng(SEED).integers(0, nnzL, K)
    bounds, s, a, q = chunk_stats(Lx)
    np.savez_compressed(
        os.path.join(OUT, "large_" + name + ".npz"),
        dims=np.array(dims), n=n, nnzL=nnzL, ref_chol_seconds=sec, seed=SEED, K=K,
        Lp_sha256=hashlib.sha256(Lp.tobytes()).hexdigest(), Li_sha256=hashlib.sha256(Li.tobytes()).hexdigest(),
        Ap_sha256=hashlib.sha256(Ap.tobytes()).hexdigest(), Ai_sha256=hashlib.sha256(Ai.tobytes()).hexdigest(),
        sample_x=Lx[pos], chunk_sum=s, chunk_abs=a, chunk_sq=q,
        diag_tail=Lx[Lp[n - 4096:n]], flops=float((cc.astype(np.float64) ** 2).sum()))
    print(f"{name}: n={n} nnzL={nnzL} ref chol {sec:.1f}s (wall {time.time() - t0:.1f}s)", flush=True)


if __name__ == "__main__":
    for nm in (sys.argv[1:] or ["lap2d_1024", "lap3d_64"]):
        make(nm)
