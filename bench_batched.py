"""bench.py leg for BASELINE config 5: nbatch SPD matrices D_b A D_b over the 1138_bus pattern, numeric
chol sharded over the ranks by batch index (no collective on the data path).  A step = one
factorization of the whole batch.  Prints the same JSON line as the single-matrix bench."""
import ctypes as C
import json
import os
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
MTX = os.path.join(ROOT, "tests", "data", "1138_bus.mtx")


def run(args, rank, world, local_rank, ClockSampler, METRIC, UNIT):
    import torch
    import sparsecholesky_b200 as sc
    from sparsecholesky_b200 import multigpu, synth
    from sparsecholesky_b200._lib import check, lib
    L = lib()
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()

    def reduce(v, op="max"):
        if dist is None:
            return float(v)
        t = torch.tensor([float(v)], dtype=torch.float64, device=f"cuda:{local_rank}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX if op == "max" else dist.ReduceOp.SUM)
        return float(t.item())

    nbatch = args.nbatch
    A = sc.load_matrix_market_to_csc(MTX)
    S = sc.schol(A)
    plan = S.plan(local_rank)
    b0, b1 = multigpu.batch_shard(nbatch, rank, world)
    nb = b1 - b0
    nA, nL = A.capacity(), S.nnz
    Ax = np.stack([synth.batch_values(A.p, A.i, A.x, b) for b in range(b0, b1)])
    dev = f"cuda:{local_rank}"
    dAx = torch.from_numpy(Ax).to(dev)
    dLx = torch.empty((nb, nL), dtype=torch.float64, device=dev)
    dinfo = torch.zeros(nb, dtype=torch.int32, device=dev)
    stream = torch.cuda.ExternalStream(L.spchol_stream(plan.h), device=local_rank)

    def step():
        check(L.spchol_factor_batched_device(plan.h, nb, dAx.data_ptr(), dLx.data_ptr(), dinfo.data_ptr()), "factor_batched")

    for _ in range(args.warmup):
        step()
    stream.synchronize()
    assert int(dinfo.abs().max().item()) == 0
    # parity spot check (the oracle as checker, outside the timed region): first, middle and last matrix of this shard
    from oracle import pyoracle as po
    parity = 0.0
    for b in sorted({0, nb // 2, nb - 1}):
        rc, _, _, ref = po.chol(A.n, A.p, A.i, Ax[b])
        assert rc == 0
        got = dLx[b].cpu().numpy()
        parity = max(parity, float((np.abs(got - ref) / np.maximum(np.abs(ref), 1e-300)).max()))
    assert parity <= 1e-10, f"batched factor differs from the oracle: {parity:.3e}"
    barrier()
    sampler = ClockSampler(local_rank); sampler.start()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    stream.synchronize()
    barrier()
    clocks = sampler.stop()
    ms = reduce(e0.elapsed_time(e1) / args.steps)
    launches = int(L.spchol_last_launches(plan.h)) * args.steps
    flops_total = S.flops * nbatch
    gflops = flops_total / ms * 1e-6
    # roofline: the whole batched level program is HBM/latency bound; algorithmic bytes per matrix
    #  = 8 nnz(A) + 16 nnz(L) (+ 8 nnz(L) for the CSC export inside the timed region)
    bytes_per_matrix = 8 * nA + 24 * nL
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    achieved = bytes_per_matrix * nb / (reduce(e0.elapsed_time(e1) / args.steps) * 1e-3) * 1e-9
    roofline = {"kernel": "k_tree_team (one 4-warp team per matrix walks the whole assembly tree in amalgamated groups; A read from Ax, L written in CSC order)", "bound": "hbm",
                "achieved": round(achieved, 1), "peak": hbm, "unit": "GB/s", "frac": round(achieved / hbm, 4), "traffic": None,
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s (of fallback)",
                "algorithmic_bytes_per_matrix": bytes_per_matrix, "matrices_per_launch_set": nb,
                "note": "per-rank; achieved = algorithmic bytes of this rank's shard / step time (SURVEY 8d: 8 nnz(A) + 24 nnz(L) per matrix)"}

    e2e = None
    if not args.no_e2e:
        hA = C.c_void_p(); hL = C.c_void_p(); hI = C.c_void_p()
        check(L.spchol_host_alloc(nb * nA * 8, C.byref(hA))); check(L.spchol_host_alloc(nb * nL * 8, C.byref(hL)))
        check(L.spchol_host_alloc(nb * 4, C.byref(hI)))
        C.memmove(hA.value, Ax.ctypes.data, nb * nA * 8)
        check(L.spchol_factor_batched(plan.h, nb, hA.value, hL.value, hI.value), "e2e warmup")
        barrier()
        k = max(1, min(args.steps, 3))
        t0 = time.perf_counter()
        for _ in range(k):
            check(L.spchol_factor_batched(plan.h, nb, hA.value, hL.value, hI.value), "e2e")
        torch.cuda.synchronize()
        e2e_ms = reduce((time.perf_counter() - t0) * 1e3 / k)
        e2e = {"value": flops_total / e2e_ms * 1e-6, "unit": UNIT, "ms_per_step": e2e_ms, "steps": k,
               "h2d_bytes_per_step": nbatch * nA * 8, "d2h_bytes_per_step": nbatch * (nL * 8 + 4),
               "result": "CSC L values of every matrix (reference layout) + info per matrix"}
        for h in (hA, hL, hI):
            L.spchol_host_free(h)

    cpu = None
    if not args.no_cpu_baseline and world == 1:
        from oracle import pyoracle as po
        nsamp = 64
        parent, Lp, _ = po.schol(A.n, A.p, A.i, want_Li=False)
        t0 = time.perf_counter()
        for b in range(nsamp):
            if po.have_ref():
                rc = po.ref_chol(A.n, A.p, A.i, Ax[b % nb], nL)[0]
            else:
                rc = po.chol(A.n, A.p, A.i, Ax[b % nb], parent, Lp)[0]
            assert rc == 0
        sec = time.perf_counter() - t0
        cpu = {"value": S.flops * nsamp / sec * 1e-9, "unit": UNIT, "cores": 1, "kind": "reference" if po.have_ref() else "port",
               "seconds": sec, "sample": f"{nsamp} of the {nbatch} matrices, chol() each (includes the per-call symbolic work the reference redoes, "
               "and the ctypes call overhead)", "host_cores_available": os.cpu_count()}

    if rank == 0:
        line = {"metric": METRIC, "value": gflops, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms, "chol_ms": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic",
                "config": {"workload": args.workload, "description": f"{nbatch} SPD matrices D_b A D_b sharing the 1138_bus pattern/S, "
                           f"batch sharded over {world} rank(s), no collective", "n": A.n, "nnz_A": nA, "nnz_L": nL, "nbatch": nbatch,
                           "flops": flops_total, "l2": f"inputs larger than L2 ({nb * (nA + nL) * 8 / 1e9:.2f} GB of values per rank per step)"},
                "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
                "parity": {"max_rel_err_vs_oracle": parity, "bound": 1e-10, "matrices_checked": len({0, nb // 2, nb - 1})}}
        print(json.dumps(line), flush=True)
    if dist is not None:
        barrier()
        dist.destroy_process_group()
