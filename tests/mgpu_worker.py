"""Worker of tests/test_multigpu.py, launched by torch.distributed.run with one rank per GPU (NCCL).
Factors a set of matrices with the multi-GPU plan and checks L on every rank against the oracle."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    import sparsecholesky_b200 as sc
    from oracle import pyoracle as po
    from sparsecholesky_b200 import multigpu, synth
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    comm = multigpu.Comm(dist, device=local)
    rank, world = comm.rank, comm.world
    gold = os.path.join(ROOT, "tests", "golden")
    cases = []
    for dims, nb, mw in (((40, 40, 1), 8, 16), ((12, 12, 12), 8, 16), ((96, 96, 1), 64, 64), ((24, 24, 24), 64, 128),
                         ((256, 256, 1), 64, 128), ((32, 32, 32), 128, 256)):
        perm = synth.nd_perm(*dims)
        ti, tj, tx, n = synth.laplacian_triplets(*dims, perm=perm)
        cases.append(("lap" + "x".join(map(str, dims)), sc.triplet_to_csc_matrix(ti, tj, tx, n), nb, mw))
    for name in ("bcsstk01", "1138_bus"):
        g = np.load(os.path.join(gold, name + ".npz"))
        cases.append((name, sc.CscMatrix(int(g["n"]), g["Ap"], g["Ai"], g["Ax"]), 4, 4))
    worst = 0.0
    worst_solve = 0.0
    for name, A, nb, mw in cases:
        S = sc.schol(A, relax=3)                  # tree-parallel amalgamation, as bench.py uses for world > 1
        plan = multigpu.DistPlan(S, comm, dist_NB=nb, dist_min_w=mw)
        st = plan.stats()
        for rep in range(2):                      # the plan is reusable
            r = multigpu.chol_dist(A, plan)
            assert r.has_value(), name
            L = r.value()
            rc, Lp, Li, Lx = po.chol(A.n, A.p, A.i, A.x)
            assert rc == 0 and (L.p == Lp).all() and (L.i == Li).all(), name
            err = float((np.abs(L.x - Lx) / np.abs(Lx)).max())
            assert err <= 1e-10, (name, err)
            worst = max(worst, err)
            # the distributed residual check (every rank multiplies the panels it holds or leads, ncclAllReduce of the probes)
            import ctypes as C
            from sparsecholesky_b200._lib import check, lib, ptr
            out = C.c_double(0)
            check(lib().spchol_residual(plan.h, ptr(np.ascontiguousarray(A.x)), C.byref(out)), "spchol_residual")
            assert out.value <= 1e-12, (name, out.value)
            # the distributed solve: every front is solved by one rank, contributions to the shared top of the tree are
            # exchanged once per level; every rank gets the whole x (1 and 11 right-hand sides: two passes of 8)
            rng = np.random.default_rng(11 + rep)
            for nrhs in (1, 11):
                import scipy.sparse as sp
                Au = sp.csc_matrix((A.x, A.i, A.p), shape=(A.n, A.n))
                Af = Au + sp.triu(Au, 1).T
                bb = Af @ rng.standard_normal((A.n, nrhs))
                xs = multigpu.solve_dist(plan, bb if nrhs > 1 else bb[:, 0]).reshape(A.n, nrhs)
                serr = float(np.linalg.norm(Af @ xs - bb) / np.linalg.norm(bb))     # solution of order one: rounding level whatever the condition number
                assert serr <= 1e-13, (name, "solve", nrhs, serr)
                worst_solve = max(worst_solve, serr)
        if rank == 0:
            print(f"mgpu {name}: world={world} dist fronts={st['fronts_dist']} transfers={st['transfers']} max rel err={err:.2e}", flush=True)
        plan.close()
    # not positive definite: every rank must report the same error
    perm = synth.nd_perm(40, 40, 1)
    ti, tj, tx, n = synth.laplacian_triplets(40, 40, 1, perm)
    tx = tx.copy(); tx[n - 1] = -1.0
    A = sc.triplet_to_csc_matrix(ti, tj, tx, n)
    plan = multigpu.DistPlan(sc.schol(A), comm, dist_NB=8, dist_min_w=16)
    r = multigpu.chol_dist(A, plan)
    assert not r.has_value() and r.error() == "A is not positive definite."
    plan.close()
    dist.barrier()
    if rank == 0:
        print(f"MGPU OK world={world} worst={worst:.2e} solve residual={worst_solve:.2e}", flush=True)
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
