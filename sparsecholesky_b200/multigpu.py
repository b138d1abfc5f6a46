"""Multi-GPU numeric Cholesky: one process per GPU, `torch.distributed` only for the rendezvous
(moving the NCCL unique id and reducing timings); the data path is libspchol's own stream-ordered
program of sm_100a kernels and NCCL transfers (include/spchol.h, "multi-GPU").

    comm = Comm(dist)                       # collective
    S = schol(A)                            # every rank analyses the same pattern
    plan = DistPlan(S, comm)                # collective: proportional mapping + sub-communicators
    L = chol_dist(A, plan)                  # collective; full CSC L on every rank (small problems)

No reference counterpart: the reference is single-threaded."""
import ctypes as C
import os
import time

import numpy as np

from ._lib import check, lib, ptr
from .api import CscMatrix, Expected, Sym


class Comm:
    """spchol_comm over the ranks of an initialised torch.distributed process group."""

    def __init__(self, dist, device=None):
        import torch
        self.dist = dist
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        self.device = int(os.environ.get("LOCAL_RANK", "0")) if device is None else device
        L = lib()
        ident = np.zeros(128, np.uint8)
        if self.rank == 0:
            check(L.spchol_comm_unique_id(ptr(ident)), "spchol_comm_unique_id")
        t = torch.from_numpy(ident)
        if dist.get_backend() == "nccl":
            t = t.to(f"cuda:{self.device}")
        dist.broadcast(t, 0)
        ident = t.cpu().numpy().copy()
        h = C.c_void_p()
        check(L.spchol_comm_create(self.rank, self.world, ptr(ident), self.device, C.byref(h)), "spchol_comm_create")
        self.h = h

    def close(self):
        if self.h:
            lib().spchol_comm_destroy(self.h)
            self.h = None


class DistPlan:
    def __init__(self, S, comm, dist_NB=0, dist_min_w=0):
        self.S, self.comm = S, comm
        opts = np.array([dist_NB, dist_min_w], np.int32)
        h = C.c_void_p()
        check(lib().spchol_plan_create_dist(S._sym.h, comm.h, ptr(opts), C.byref(h)), "spchol_plan_create_dist")
        self.h = h

    def stats(self):
        st = np.zeros(10, np.int64); ds = np.zeros(8, np.int64); load = np.zeros(self.comm.world)
        check(lib().spchol_plan_stats(self.h, ptr(st)))
        check(lib().spchol_plan_dist_stats(self.h, ptr(ds), ptr(load)))
        keys = ["ns", "nlevels", "n_small", "n_large", "lbuf_entries", "ws_entries", "n_ops", "max_m", "max_w", "device_bytes"]
        d = dict(zip(keys, (int(v) for v in st)))
        d.update(dict(zip(["rank", "world", "fronts_held", "fronts_dist", "transfers", "doubles_moved", "dist_NB", "readback_entries"], (int(v) for v in ds[:8]))))
        d["rank_load"] = load.tolist()
        return d

    def close(self):
        if self.h:
            lib().spchol_plan_destroy(self.h)
            self.h = None


def map_fronts(S, world, dist_min_w=0):
    """Host-only: the rank groups a `world`-rank plan gives every (relaxed) front."""
    ns = S.stats()["ns_relaxed"]
    lo = np.empty(ns, np.int32); cnt = np.empty(ns, np.int32); load = np.zeros(world)
    check(lib().spchol_map_fronts(S._sym.h, world, dist_min_w, ptr(lo), ptr(cnt), ptr(load)))
    return lo, cnt, load


def plan_dryrun(S, rank, world, dist_NB=0, dist_min_w=0):
    """Host-only planner statistics of one rank (no GPU, no NCCL)."""
    st = np.zeros(10, np.int64); ex = np.zeros(8, np.int64); sig = np.zeros(4, np.int64)
    opts = np.array([dist_NB, dist_min_w], np.int32)
    check(lib().spchol_plan_dryrun_dist(S._sym.h, rank, world, ptr(opts), ptr(st), ptr(ex), ptr(sig)))
    return {"lbuf_entries": int(st[4]), "ws_entries": int(st[5]), "n_ops": int(st[6]), "device_bytes": int(st[9]),
            "gemm_flops": int(ex[5]), "transfers": int(ex[6]), "doubles_moved": int(ex[7]),
            "bcasts": int(sig[0]), "sends": int(sig[1]), "recvs": int(sig[2]), "comm_hash": int(sig[3])}


def chol_dist(A, plan):
    """Collective numeric factorization; returns Expected[CscMatrix] with the full L on every rank
    (each rank exports the entries it holds, the vectors are summed)."""
    import torch
    S, comm = plan.S, plan.comm
    Lx = np.empty(S.nnz, np.float64)
    info = C.c_int(0)
    Ax = np.ascontiguousarray(A.x, np.float64)
    check(lib().spchol_factor(plan.h, ptr(Ax), ptr(Lx), C.byref(info)), "chol_dist")
    t = torch.from_numpy(Lx)
    if comm.dist.get_backend() == "nccl":
        t = t.to(f"cuda:{comm.device}")
    comm.dist.all_reduce(t)
    Lx = t.cpu().numpy()
    if info.value > 0:
        return Expected(error="A is not positive definite.")
    return Expected(value=CscMatrix(A.n, S.p, S.i, Lx, Sym.none))


def solve_dist(plan, b):
    """Collective x = A^-1 b with the factor the plan holds after chol_dist (spchol_solve on a multi-rank plan): every rank
    passes the same b (n,) or (n, nrhs) and receives the whole x.  Each front is solved by the rank that holds or leads it;
    the ranks exchange their contributions to the shared top of the tree once per level (ncclAllReduce)."""
    B = np.asfortranarray(np.asarray(b, np.float64).reshape(plan.S.n, -1))
    X = np.empty_like(B, order="F")
    check(lib().spchol_solve(plan.h, B.ctypes.data, X.ctypes.data, B.shape[1]), "solve_dist")
    return X.reshape(np.shape(b))


def batch_shard(nbatch, rank, world):
    """[b0, b1) of the batch index this rank factors (config 5: no collective on the data path)."""
    base, rem = divmod(nbatch, world)
    b0 = rank * base + min(rank, rem)
    return b0, b0 + base + (1 if rank < rem else 0)
