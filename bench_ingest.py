"""Ingest timings for bench.py's `also.ingest`: COO -> CSC and the Matrix Market parse on the GPU against the
product's host paths on the same inputs, bit-exact compare included.  One JSON line.  `python bench_ingest.py`"""
import ctypes as C
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def main():
    import sparsecholesky_b200 as sc
    from sparsecholesky_b200 import synth
    from sparsecholesky_b200._lib import check, lib, ptr
    L = lib()
    out = {}
    # ---- COO -> CSC, config 3's matrix: 2D 1024^2, nested-dissection order, 3.1 M triplets
    dims = (1024, 1024, 1)
    ti, tj, tx, n = synth.laplacian_triplets(*dims, perm=synth.nd_perm(*dims))
    t0 = time.perf_counter()
    H = sc.triplet_to_csc_matrix(ti, tj, tx, n)
    host_s = time.perf_counter() - t0
    sc.triplet_to_csc_matrix_gpu(ti, tj, tx, n)                                 # work space, kernels loaded
    ms = min(sc.triplet_to_csc_matrix_gpu(ti, tj, tx, n, return_ms=True)[1] for _ in range(3))
    A = sc.triplet_to_csc_matrix_gpu(ti, tj, tx, n)
    out["coo_to_csc"] = {"matrix": "lap2d_1024", "triplets": int(len(ti)), "device_ms": round(ms, 3), "host_ms": round(host_s * 1e3, 1),
                         "bit_exact_vs_host": bool(np.array_equal(A.p, H.p) and np.array_equal(A.i, H.i) and np.array_equal(A.x, H.x)),
                         "entry": "spchol_triplet_to_csc_device (timed inside spchol_triplet_to_csc_gpu with CUDA events, inputs resident)"}
    # ---- Matrix Market text -> triplets: 2D 512^2 with random values as a 25.7 MB coordinate file
    dims = (512, 512, 1)
    ti, tj, tx, n = synth.laplacian_triplets(*dims, perm=synth.nd_perm(*dims))
    rng = np.random.default_rng(3)
    tx = tx * (1.0 + rng.random(len(tx)))
    with tempfile.TemporaryDirectory() as tmp:
        path = os.path.join(tmp, "lap2d_512.mtx")
        with open(path, "w") as fh:
            fh.write("%%MatrixMarket matrix coordinate real symmetric\n%\n")
            fh.write(f"{n} {n} {len(ti)}\n")
            fh.write("".join("%d %d %.17g\n" % t for t in zip((ti + 1).tolist(), (tj + 1).tolist(), tx.tolist())))
        size = os.path.getsize(path)
        nn = C.c_int(0); nz = C.c_int64(0); dms = C.c_double(0)
        check(L.spchol_mtx_read(path.encode(), C.byref(nn), C.byref(nz), None, None, None), "mtx_read")
        hi = np.empty(nz.value, np.int32); hj = np.empty(nz.value, np.int32); hx = np.empty(nz.value)
        t0 = time.perf_counter()
        check(L.spchol_mtx_read(path.encode(), C.byref(nn), C.byref(nz), ptr(hi), ptr(hj), ptr(hx)), "mtx_read")
        host_s = time.perf_counter() - t0
        gi = np.empty(nz.value, np.int32); gj = np.empty(nz.value, np.int32); gx = np.empty(nz.value)
        best = 1e30
        for _ in range(3):
            check(L.spchol_mtx_read_gpu(path.encode(), C.byref(nn), C.byref(nz), ptr(gi), ptr(gj), ptr(gx), C.byref(dms)), "mtx_read_gpu")
            best = min(best, dms.value)
    out["mtx_parse"] = {"matrix": "lap2d_512, %.17g values", "file_bytes": int(size), "entries": int(nz.value), "device_ms": round(best, 3),
                        "device_gb_per_s": round(size / best * 1e-6, 2), "host_ms": round(host_s * 1e3, 1),
                        "bit_exact_vs_host": bool(np.array_equal(gi, hi) and np.array_equal(gj, hj) and np.array_equal(gx.view(np.uint64), hx.view(np.uint64))),
                        "entry": "spchol_mtx_parse_device (timed inside spchol_mtx_read_gpu with CUDA events, text resident)"}
    assert out["coo_to_csc"]["bit_exact_vs_host"] and out["mtx_parse"]["bit_exact_vs_host"]
    print(json.dumps(out))


if __name__ == "__main__":
    main()
