"""Quick device timing of chol(A,S) on a few Laplacians (development aid, not the bench)."""
import ctypes as C
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import sparsecholesky_b200 as sc
from sparsecholesky_b200 import synth
from sparsecholesky_b200._lib import lib, ptr

cases = [(256, 256, 1), (512, 512, 1), (1024, 1024, 1), (32, 32, 32), (48, 48, 48), (64, 64, 64)]
if len(sys.argv) > 1:
    cases = [tuple(int(v) for v in a.split("x")) for a in sys.argv[1:]]
for dims in cases:
    t0 = time.time()
    perm = synth.nd_perm(*dims)
    ti, tj, tx, n = synth.laplacian_triplets(*dims, perm=perm)
    A = sc.triplet_to_csc_matrix(ti, tj, tx, n)
    t1 = time.time()
    S = sc.schol(A)
    t2 = time.time()
    plan = S.plan(0)
    t3 = time.time()
    L = sc.chol(A, S)
    ok = L.has_value()
    best = None
    for _ in range(3):
        r = sc.chol(A, S)
        ms = np.zeros(4)
        lib().spchol_last_timing(plan.h, ptr(ms))
        if best is None or ms[1] < best[1]:
            best = ms.copy()
    lib().spchol_set_profile(plan.h, 1)
    sc.chol(A, S) if S.nnz < 2**31 else None
    kms = np.zeros(12); kl = np.zeros(12, np.int64); fl2 = np.zeros(2)
    lib().spchol_get_profile(plan.h, ptr(kms), ptr(kl), ptr(fl2))
    kms[5] = kms[9:12].sum(); kl[5] = kl[9:12].sum()      # DMMA updates are timed in three K classes
    lib().spchol_set_profile(plan.h, 0)
    names = ["small", "zero", "extadd", "potrf", "trsm", "gemm", "memset+asm"]
    print("   per-kind ms: " + " ".join(f"{nm}={kms[i]:.2f}({kl[i]})" for i, nm in enumerate(names)) +
          f"  gemm alg TF/s={fl2[0]/max(kms[5],1e-9)*1e-9:.2f} (alg/full flops {fl2[0]/max(fl2[1],1):.2f})")
    st = S.stats(); ps = S.plan_stats()
    print(f"{dims}: n={n} nnzL={S.nnz} flops={S.flops:.3e} ok={ok} gen={t1-t0:.2f}s schol={t2-t1:.2f}s plan={t3-t2:.2f}s "
          f"h2d={best[0]:.2f} numeric={best[1]:.2f} export={best[2]:.2f} d2h={best[3]:.2f} ms  "
          f"GF/s={S.flops/best[1]*1e-6:.1f} launches={lib().spchol_last_launches(plan.h)} "
          f"fronts={ps['ns']} small={ps['n_small']} large={ps['n_large']} levels={ps['nlevels']} "
          f"lbuf={ps['lbuf_entries']*8/1e9:.2f}GB ws={ps['ws_entries']*8/1e9:.2f}GB", flush=True)
    del L, S, plan
