#!/usr/bin/env python
"""bench.py -- fp64 factor GFLOP/s + chol() ms of the B200 numeric sparse Cholesky (BASELINE.json metric).

    python bench.py --gpus 1 --steps K --warmup W [--workload lap3d_128]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...        # the reference's own CPU chol() on a bounded sample

A step = one numeric factorization chol(A, S) of the workload with the symbolic plan resident
(the plan is what `S` carries between calls).  `value` = sum_j colcount_j^2 / device time with the
values of A already in HBM; `e2e` = the same through the host-buffer C-ABI call (pinned host A in,
L back to pinned host memory inside the timed region).  Prints ONE JSON line on rank 0.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (dims, description, cpu sample dims)
    "lap3d_128": ((128, 128, 128), "3D 7-point Laplacian 128^3 (2.1M rows), geometric ND", (32, 32, 32)),
    "lap3d_96": ((96, 96, 96), "3D 7-point Laplacian 96^3, geometric ND", (32, 32, 32)),
    "lap3d_64": ((64, 64, 64), "3D 7-point Laplacian 64^3, geometric ND", (32, 32, 32)),
    "lap3d_32": ((32, 32, 32), "3D 7-point Laplacian 32^3, geometric ND", (32, 32, 32)),
    "lap2d_1024": ((1024, 1024, 1), "2D 5-point Laplacian 1024^2 (1.05M rows), geometric ND", (256, 256, 1)),
    "lap2d_256": ((256, 256, 1), "2D 5-point Laplacian 256^2, geometric ND", (256, 256, 1)),
    # config 5: handled by bench_batched.py (next to this file)
    "batched_1138": (None, "16k SPD matrices over the 1138_bus pattern (batched numeric chol)", None),
}
METRIC = "fp64_factor_gflops"
UNIT = "GFLOP/s"


def make_matrix(dims):
    import sparsecholesky_b200 as sc
    from sparsecholesky_b200 import synth
    perm = synth.nd_perm(*dims)
    ti, tj, tx, n = synth.laplacian_triplets(*dims, perm=perm)
    return sc.triplet_to_csc_matrix(ti, tj, tx, n)


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self._stop, self._t = index, [], threading.Event(), None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([v.strip() for v in out.strip().split(",")])
            except Exception:
                pass
            self._stop.wait(0.1)

    def start(self):
        self._t = threading.Thread(target=self._run, daemon=True); self._t.start()

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join(timeout=6)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except ValueError:
                continue
            for k, nm in enumerate(names):
                if r[3 + k].lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_reference_time(dims, repeats=1):
    """The reference's own chol() (oracle/_ref/libref.so when it was built from /root/reference,
    else the plain-C port) on one host core.  Returns (kind, seconds_best, flops, n, nnzL)."""
    from oracle import pyoracle as po
    A = make_matrix(dims)
    n = A.n
    parent, Lp, _ = po.schol(n, A.p, A.i, want_Li=False)
    nnzL = int(Lp[n])
    cc = np.diff(Lp).astype(np.float64)
    flops = float((cc * cc).sum())
    best = None
    if po.have_ref():
        kind = "reference"
        for _ in range(repeats):
            rc, _, _, _, sec, err = po.ref_chol(n, A.p, A.i, A.x, nnzL)
            assert rc == 0, err
            best = sec if best is None else min(best, sec)
    else:
        kind = "port"
        for _ in range(repeats):
            t0 = time.perf_counter()
            rc, _, _, _ = po.chol(n, A.p, A.i, A.x, parent, Lp)
            sec = time.perf_counter() - t0
            assert rc == 0
            best = sec if best is None else min(best, sec)
    return kind, best, flops, n, nnzL


def make_matrix_checker_only(dims):
    """The same matrix as make_matrix() built WITHOUT the product library (numpy ND permutation + the checker's own
    triplet_to_csc): the reference arm maps only oracle/ into its process.  -> (n, Ap, Ai, Ax)"""
    from oracle import pyoracle as po
    from sparsecholesky_b200 import synth          # pure numpy helpers; the package does not load libspchol.so on import
    perm = synth.nd_perm_py(*dims)
    ti, tj, tx, n = synth.laplacian_triplets(*dims, perm=perm)
    tri = po.ref_triplet_to_csc if po.have_ref() else po.triplet_to_csc
    Ap, Ai, Ax = tri(ti, tj, tx, n)
    return n, Ap, Ai, Ax


def run_reference(args, rank):
    if rank != 0:
        return
    dims, desc, sample = WORKLOADS[args.workload]
    from oracle import pyoracle as po
    n, Ap, Ai, Ax = make_matrix_checker_only(sample)

    class _A:
        pass
    A = _A(); A.n, A.p, A.i, A.x = n, Ap, Ai, Ax
    parent, Lp, _ = po.schol(n, A.p, A.i, want_Li=False)
    nnzL = int(Lp[n])
    cc = np.diff(Lp).astype(np.float64)
    flops = float((cc * cc).sum())
    use_ref = po.have_ref()

    def step():
        if use_ref:
            rc, _, _, _, sec, err = po.ref_chol(n, A.p, A.i, A.x, nnzL)
            assert rc == 0, err
            return sec
        t0 = time.perf_counter()
        rc, _, _, _ = po.chol(n, A.p, A.i, A.x, parent, Lp)
        assert rc == 0
        return time.perf_counter() - t0

    for _ in range(args.warmup):
        step()
    t = [step() for _ in range(args.steps)]
    sec = float(np.mean(t))
    gf = flops / sec * 1e-9
    smp = f"{'x'.join(map(str, sample))} Laplacian (same generator/ND ordering), chol() only, symbolic excluded"
    line = {
        "impl": "reference", "metric": METRIC, "value": gf, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": args.workload, "description": desc, "reference_sample": smp,
                   "note": "the reference cannot run 128^3 (32-bit nnz overflow) and needs ~10 min for 64^3; "
                           "its GFLOP/s falls with n (O(n^2) workspace), so this sample flatters it.  The SAME-INPUT "
                           "comparison is in the product arm's line: also.lap3d_32.same_input_speedup (the 32^3 matrix "
                           "this arm times, factored by the GPU path through the host-buffer call)"},
        "seconds_per_step": sec,
        "cpu_baseline": {"value": gf, "unit": UNIT, "cores": 1, "kind": "reference" if use_ref else "port", "sample": smp},
        "e2e": {"value": gf, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def run_also():
    """lap2d_1024 (config 3), batched_1138 (config 5), lap3d_32 (the matrix the reference arm times): one child process
    each, same code path as a direct `bench.py --workload ...` run; returns {workload: summary}."""
    out = {}
    jobs = [("lap2d_1024", ["--steps", "10", "--warmup", "3", "--no-cpu-baseline"]),
            ("batched_1138", ["--steps", "5", "--warmup", "3", "--no-cpu-baseline"]),
            ("lap3d_32", ["--steps", "10", "--warmup", "3"])]
    for wl, extra in jobs:
        t0 = time.time()
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--workload", wl, "--no-also"] + extra,
                               capture_output=True, text=True, timeout=600)
            line = next((l for l in r.stdout.splitlines() if l.startswith("{")), None)
            if r.returncode != 0 or line is None:
                out[wl] = {"error": (r.stderr or r.stdout)[-300:]}
                continue
            d = json.loads(line)
            e = d.get("e2e") or {}
            rf = d.get("roofline") or {}
            sm = {"ms_per_step": d["ms_per_step"], "value": d["value"], "unit": d["unit"], "steps": d["steps"], "gpu_launches": d["gpu_launches"],
                  "e2e": {k: e.get(k) for k in ("value", "ms_per_step", "h2d_bytes_per_step", "d2h_bytes_per_step", "result")},
                  "roofline": {k: rf.get(k) for k in ("kernel", "bound", "achieved", "peak", "unit", "frac", "whole_factor_frac_of_peak")},
                  "kernel_ms": d.get("kernel_ms"), "residual": (d.get("residual") or {}).get("rel_fro_estimate"),
                  "parity": d.get("parity"), "clocks": d.get("clocks"), "wall_seconds": round(time.time() - t0, 1)}
            cpu = d.get("cpu_baseline")
            if wl == "lap3d_32" and cpu and e.get("ms_per_step"):
                sm["cpu_baseline"] = cpu
                sm["same_input_speedup"] = cpu["seconds"] / (e["ms_per_step"] * 1e-3)
                sm["same_input_note"] = ("reference chol() seconds on this matrix / seconds of the GPU path through the host-buffer "
                                         "C-ABI call (H2D + factor + CSC export + D2H), same process, same input")
            out[wl] = sm
        except Exception as exc:
            out[wl] = {"error": repr(exc)[:300]}
    # the ingest in front of the path: COO -> CSC and the Matrix Market parse on the device, bench_ingest.py
    t0 = time.time()
    try:
        r = subprocess.run([sys.executable, os.path.join(os.path.dirname(os.path.abspath(__file__)), "bench_ingest.py")],
                           capture_output=True, text=True, timeout=300)
        line = next((l for l in r.stdout.splitlines() if l.startswith("{")), None)
        out["ingest"] = json.loads(line) if r.returncode == 0 and line else {"error": (r.stderr or r.stdout)[-300:]}
        out["ingest"]["wall_seconds"] = round(time.time() - t0, 1)
    except Exception as exc:
        out["ingest"] = {"error": repr(exc)[:300]}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="lap3d_128", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-also", action="store_true", help="default workload only: skip the extra lines for configs 3, 5 and the same-input 32^3 case")
    ap.add_argument("--nbatch", type=int, default=16384, help="batched_1138: matrices in the batch (whole job)")
    ap.add_argument("--dist-nb", type=int, default=0, help="block-column width of distributed fronts (0: library default)")
    ap.add_argument("--dist-min-w", type=int, default=0, help="narrowest front factored by a rank group (0: default)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        if args.workload == "batched_1138":
            args.workload = "lap3d_128"          # the reference arm times the headline workload's bounded sample
        run_reference(args, rank)
        return
    if args.workload == "batched_1138":
        import bench_batched
        bench_batched.run(args, rank, world, local_rank, ClockSampler, METRIC, UNIT)
        return

    import torch
    import sparsecholesky_b200 as sc
    from sparsecholesky_b200._lib import check, lib, ptr
    L = lib()
    if sc.device_count() <= 0:
        raise SystemExit("bench.py: no CUDA device (the numeric path has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dist = None
    comm = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        from sparsecholesky_b200 import multigpu
        comm = multigpu.Comm(dist, device=local_rank)

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()

    def max_over_ranks(v):
        if dist is None:
            return float(v)
        t = torch.tensor([float(v)], dtype=torch.float64, device=f"cuda:{local_rank}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def gather_list(v):
        if dist is None:
            return [v]
        out = [None] * world
        dist.all_gather_object(out, v)
        return out

    dims, desc, sample = WORKLOADS[args.workload]
    t0 = time.time()
    A = make_matrix(dims)
    t_gen = time.time() - t0
    t0 = time.time()
    S = sc.schol(A, relax=3 if world > 1 else 1)      # relax 3: tree-parallel amalgamation (see spchol.h)
    t_sym = time.time() - t0
    t0 = time.time()
    if comm is not None:
        dplan = multigpu.DistPlan(S, comm, dist_NB=args.dist_nb, dist_min_w=args.dist_min_w)
        plan_h = dplan.h
        ps = dplan.stats()
    else:
        plan = S.plan(local_rank)
        plan_h = plan.h
        ps = S.plan_stats(local_rank)
    t_plan = time.time() - t0
    stream = torch.cuda.ExternalStream(L.spchol_stream(plan_h), device=local_rank)
    dAx = torch.from_numpy(A.x).to(f"cuda:{local_rank}")
    info = C.c_int(0)

    def step():
        check(L.spchol_factor_device(plan_h, dAx.data_ptr()), "spchol_factor_device")

    for _ in range(args.warmup):
        step()
    check(L.spchol_sync(plan_h, C.byref(info)))
    assert info.value == 0, f"factorization failed: info={info.value}"
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    check(L.spchol_sync(plan_h, C.byref(info)))
    barrier()
    clocks = sampler.stop()
    ms_rank = e0.elapsed_time(e1) / args.steps
    ms = max_over_ranks(ms_rank)                       # device time, max over ranks
    launches = int(L.spchol_last_launches(plan_h)) * args.steps
    gflops = S.flops / ms * 1e-6

    # roofline of the dominant kernel (DMMA SYRK/GEMM): one extra profiled step, CUDA events per launch group
    check(L.spchol_set_profile(plan_h, 1))
    step()
    check(L.spchol_sync(plan_h, C.byref(info)))
    kms = np.zeros(12); kl = np.zeros(12, np.int64); fl2 = np.zeros(2)
    check(L.spchol_get_profile(plan_h, ptr(kms), ptr(kl), ptr(fl2)))
    check(L.spchol_set_profile(plan_h, 0))
    barrier()
    peak = C.c_double(0)
    check(L.spchol_fp64_peak(local_rank, C.byref(peak)))
    fl3 = np.zeros(3)
    check(L.spchol_get_gemm_classes(plan_h, ptr(fl3)))
    names = ["small_fronts", "zero", "extend_add", "potrf", "trsm", "gemm_dmma", "transfers_nccl", "pack_unpack", "memset_assemble"]
    # the DMMA update launches are timed in three K classes (profile slots 9..11)
    kms[5] = kms[9:12].sum(); kl[5] = kl[9:12].sum()
    kind_ms = {nm: round(float(kms[i]), 3) for i, nm in enumerate(names)}
    gemm_classes = {nm: {"ms": round(float(kms[9 + i]), 3), "launches": int(kl[9 + i]),
                         "tflops": round(float(fl3[i] / kms[9 + i] * 1e-9), 2) if kms[9 + i] > 0 else None}
                    for i, nm in enumerate(["K<=128", "K<=256", "K>256"])}
    gemm_ms = float(kms[5]); gemm_launches = int(kl[5])
    achieved = fl2[0] / gemm_ms * 1e-9 if gemm_ms > 0 else 0.0
    # DRAM traffic of the dominant kernel per launch: from the committed `ncu --set full` capture of this command
    # (profiles/ncu_gemm_traffic.json, written by tools/ncu_traffic.py), not measured live
    traffic, traffic_src = None, None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "ncu_gemm_traffic.json")))
        if tj.get("workload") == args.workload:
            traffic, traffic_src = tj["dram_bytes_per_launch"], tj["source"]
    except Exception:
        pass
    gb = C.c_double(0); gt = C.c_int64(0)
    check(L.spchol_plan_gemm_bytes(plan_h, C.byref(gb), C.byref(gt)))
    kname = "k_gemm_dmma" if os.environ.get("SPCHOL_GEMM_TMA") == "0" else "k_gemm_tma"
    roofline = {
        "kernel": kname, "bound": "tensor", "achieved": round(achieved, 3), "peak": round(peak.value, 3),
        "unit": "TFLOP/s", "frac": round(achieved / peak.value, 4) if peak.value else None, "traffic": traffic,
        "traffic_source": traffic_src,
        "operand_bytes_per_launch_if_moved_once": round(gb.value / max(gemm_launches, 1), 1),
        "peak_source": "measured live: register-resident mma.sync.m8n8k4.f64 chains on this GPU (MEASURED_PEAKS.json has no FP64 entry: fallback)",
        "launches": gemm_launches, "avg_launch_ms": round(gemm_ms / max(gemm_launches, 1), 4),
        "algorithmic_flops_per_step": fl2[0], "share_of_step": round(gemm_ms / max(float(kms[:9].sum()), 1e-9), 3),
        "by_K_class": gemm_classes,
        "whole_factor_frac_of_peak": round(gflops * 1e-3 / (peak.value * world), 4) if peak.value else None,
        "note": "per-rank figures of rank 0; whole_factor_frac_of_peak = value / (n_gpus * peak)" if world > 1 else None,
    }
    per_rank = gather_list({"ms": round(ms_rank, 3), "kernel_ms": kind_ms, "gemm_tflops": round(achieved, 2),
                            "lbuf_gb": round(ps["lbuf_entries"] * 8e-9, 2), "ws_gb": round(ps["ws_entries"] * 8e-9, 2)})

    # e2e: host-buffer C-ABI call, pinned host memory, H2D + D2H inside the timed region
    e2e = None
    if not args.no_e2e:
        nA = A.capacity()
        use_csc = S.nnz <= 2**31 - 1 and world == 1
        out_entries = S.nnz if use_csc else ps["lbuf_entries"]
        hA = C.c_void_p(); hL = C.c_void_p()
        check(L.spchol_host_alloc(nA * 8, C.byref(hA)), "host_alloc")
        check(L.spchol_host_alloc(out_entries * 8, C.byref(hL)), "host_alloc")
        C.memmove(hA.value, A.x.ctypes.data, nA * 8)
        fn = L.spchol_factor if use_csc else L.spchol_factor_lbuf
        e2e_steps = max(1, min(args.steps, 3))
        check(fn(plan_h, hA.value, hL.value, C.byref(info)), "e2e warmup")
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            check(fn(plan_h, hA.value, hL.value, C.byref(info)), "e2e")
        torch.cuda.synchronize()
        e2e_ms = max_over_ranks((time.perf_counter() - t0) * 1e3 / e2e_steps)
        barrier()
        tm = np.zeros(4)
        L.spchol_last_timing(plan_h, ptr(tm))
        e2e = {"value": S.flops / e2e_ms * 1e-6, "unit": UNIT, "ms_per_step": e2e_ms, "steps": e2e_steps,
               "h2d_bytes_per_step": nA * 8 * world,
               "d2h_bytes_per_step": sum(gather_list((ps["readback_entries"] if world > 1 else out_entries) * 8 + 4)),
               "result": "CSC L values (reference layout)" if use_csc else
                         ("supernodal L panels, each rank reads back the fronts it holds alone and the block columns it owns of group fronts" if world > 1 else
                          "supernodal L panels (nnz(L) > INT_MAX: no 32-bit CSC)"),
               "last_step_ms": {"h2d": tm[0], "numeric": tm[1], "export": tm[2], "d2h": tm[3]}}
        L.spchol_host_free(hA)
        e2e_host_L = (hL, out_entries, use_csc)          # checked against the device copy below, then freed

    # full-size parity property: ||A - L L^T||_F / ||A||_F of the factor just computed (device estimate, 8 probes)
    resid = None
    if world == 1:
        step()
        check(L.spchol_sync(plan_h, C.byref(info)))
        # the panels the e2e call delivered to the host are the panels on the device (the factorization is deterministic,
        # so the device copy of this step must equal them bit for bit): 64 chunks of 1 Mi doubles spread over the buffer
        if e2e is not None and not e2e_host_L[2]:
            try:
                hL_, ne_, _ = e2e_host_L
                dptr = C.c_void_p(); dn = C.c_int64(0)
                check(L.spchol_lbuf(plan_h, C.byref(dptr), C.byref(dn)))

                class _Dev:
                    __cuda_array_interface__ = {"shape": (int(dn.value),), "typestr": "<f8", "data": (int(dptr.value), False), "version": 2}
                dt = torch.as_tensor(_Dev(), device=f"cuda:{local_rank}")
                host = np.ctypeslib.as_array(C.cast(hL_, C.POINTER(C.c_double)), shape=(ne_,))
                chunk = 1 << 20
                starts = np.linspace(0, max(ne_ - chunk, 0), 64).astype(np.int64)
                bad = sum(int(not np.array_equal(dt[a:a + chunk].cpu().numpy(), host[a:a + chunk], equal_nan=True)) for a in starts)
                e2e["readback_check"] = {"chunks": len(starts), "doubles_per_chunk": chunk, "mismatching_chunks": bad}
            except Exception as exc:                      # a diagnostic, never the reason a bench line is lost
                e2e["readback_check"] = {"error": repr(exc)[:200]}
    if e2e is not None:
        L.spchol_host_free(e2e_host_L[0])
    if True:
        # every rank calls (collective for a multi-GPU plan: each rank multiplies the panels it holds or leads, the probe
        # products are all-reduced); the factor checked is the one the timed steps produced
        out = C.c_double(0)
        barrier()
        t0 = time.perf_counter()
        check(L.spchol_residual(plan_h, ptr(np.ascontiguousarray(A.x)), C.byref(out)), "spchol_residual")
        resid = {"rel_fro_estimate": out.value, "bound": 1e-12, "probes": 8, "seconds": round(time.perf_counter() - t0, 3),
                 "ranks": world,
                 "how": "sqrt(mean_k ||(A - L L^T) z_k||^2) / ||A||_F on the device-resident supernodal factor (spchol_residual)"
                        + ("; panels multiplied by the rank that holds or leads each front, probe products summed with ncclAllReduce" if world > 1 else "")}
        assert out.value <= 1e-12, f"residual {out.value} exceeds the 1e-12 bound"

    # consumer of the factor: x = A^-1 b on the device-resident panels (spchol_solve), checked end to end
    # (multi-rank plans: collective, every rank passes the same b and gets the whole x)
    solve = None
    if True:
        import scipy.sparse as sp
        Au = sp.csc_matrix((A.x, A.i, A.p), shape=(A.n, A.n))
        Af = Au + sp.triu(Au, 1).T
        xt = np.random.default_rng(1).standard_normal(A.n)
        bv = np.asfortranarray((Af @ xt).reshape(A.n, 1))
        xv = np.empty_like(bv, order="F")
        check(L.spchol_solve(plan_h, bv.ctypes.data, xv.ctypes.data, 1), "spchol_solve")     # builds the task lists
        t0 = time.perf_counter()
        check(L.spchol_solve(plan_h, bv.ctypes.data, xv.ctypes.data, 1), "spchol_solve")
        sec = time.perf_counter() - t0
        solve = {"seconds_host_to_host": round(sec, 4), "nrhs": 1, "launches": int(L.spchol_last_launches(plan_h)), "ranks": world,
                 "rel_residual": float(np.linalg.norm(Af @ xv[:, 0] - bv[:, 0]) / np.linalg.norm(bv)),
                 "max_abs_error_vs_known_x": float(np.abs(xv[:, 0] - xt).max())}

    cpu = None
    if not args.no_cpu_baseline and world == 1:
        kind, sec, fl, n_s, nnz_s = cpu_reference_time(sample)
        cpu = {"value": fl / sec * 1e-9, "unit": UNIT, "cores": 1, "kind": kind, "seconds": sec,
               "sample": f"{'x'.join(map(str, sample))} Laplacian, same generator + ND ordering (n={n_s}, nnz(L)={nnz_s}); "
                         f"chol() only; the reference is single-threaded and cannot run the full workload "
                         f"({'32-bit nnz overflow' if S.nnz > 2**31 - 1 else 'O(n^2) workspace: minutes to hours'})",
               "host_cores_available": os.cpu_count()}

    # BASELINE configs 3 and 5 and the same-input case of the reference arm, each a short run of this same script in a
    # child process AFTER the timed region (so the driver-run line carries them); compact summaries only
    also = None
    if world == 1 and rank == 0 and args.workload == "lap3d_128" and not args.no_also:
        also = run_also()

    if rank == 0:
        cfg = {"workload": args.workload, "description": desc, "n": A.n, "nnz_A": A.capacity(), "nnz_L": S.nnz,
               "flops": S.flops, "fronts": ps["ns"], "levels": ps["nlevels"], "l2": "inputs larger than L2 "
               f"(L buffer {ps['lbuf_entries'] * 8 / 1e9:.2f} GB + workspace {ps['ws_entries'] * 8 / 1e9:.2f} GB rewritten every step)",
               "host_seconds": {"generate": round(t_gen, 2), "schol": round(t_sym, 2), "plan_upload": round(t_plan, 2)}}
        if world > 1:
            cfg["parallelism"] = (f"etree subtrees -> {world} ranks (proportional mapping); {ps['fronts_dist']} top fronts factored by rank "
                                  f"groups, block-column cyclic (NB={ps['dist_NB']}), NCCL broadcast/send/recv in stream order")
            cfg["rank_load_flops"] = ps["rank_load"]
        line = {
            "metric": METRIC, "value": gflops, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "chol_ms": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": cfg,
            "roofline": roofline, "kernel_ms": kind_ms, "cpu_baseline": cpu, "e2e": e2e, "residual": resid, "solve": solve,
            "gpu_launches": launches,
            "clocks": clocks,
        }
        if world > 1:
            line["per_rank"] = per_rank
        if also is not None:
            line["also"] = also
        print(json.dumps(line), flush=True)
    if dist is not None:
        barrier()
        dplan.close(); comm.close()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
