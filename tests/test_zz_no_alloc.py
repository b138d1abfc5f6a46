"""spchol_triplet_to_csc_device_noalloc: the device ingest with its work space in a static device array (no cudaMalloc),
for callers under CUDA graph capture or in serving loops.  Called through the C-ABI and compared bit for bit with the
existing entry point spchol_triplet_to_csc_device, two calls in a row (the second on another input: the shared array
keeps the first call's contents).  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from sparsecholesky_b200._lib import check, lib

pytestmark = pytest.mark.gpu

MAX_NZ = 1048576          # SPCHOL_NOALLOC_MAX_NZ / SPCHOL_NOALLOC_MAX_N of include/spchol.h
MAX_N = 1048576


def _triplets(n, nz, seed):
    rng = np.random.default_rng(seed)
    ti = rng.integers(0, n, nz).astype(np.int32)
    tj = rng.integers(0, n, nz).astype(np.int32)
    tx = rng.standard_normal(nz)
    tx[rng.integers(0, nz, max(1, nz // 10))] = 0.0          # explicit zeros stay structural
    return ti, tj, tx


def _run(fn, dev, n, ti, tj, tx):
    """one call of an ingest entry point on device buffers (torch: device memory only); returns host Ap, Ai, Ax"""
    import torch
    nz = len(ti)
    d_ti = torch.from_numpy(ti).to(dev); d_tj = torch.from_numpy(tj).to(dev); d_tx = torch.from_numpy(tx).to(dev)
    d_Ap = torch.full((n + 1,), -7, dtype=torch.int32, device=dev)
    d_Ai = torch.full((max(nz, 1),), -7, dtype=torch.int32, device=dev)
    d_Ax = torch.full((max(nz, 1),), float("nan"), dtype=torch.float64, device=dev)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    nnz = C.c_int64(-1)
    rc = fn(nz, d_ti.data_ptr(), d_tj.data_ptr(), d_tx.data_ptr(), n, d_Ap.data_ptr(), d_Ai.data_ptr(), d_Ax.data_ptr(), C.byref(nnz), None)
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    check(rc, "ingest")
    k = nnz.value
    return d_Ap.cpu().numpy(), d_Ai[:k].cpu().numpy(), d_Ax[:k].cpu().numpy(), free0 - free1


def test_noalloc_ingest_matches_the_allocating_entry_point_twice_in_a_row():
    import torch
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    L = lib()
    # first a large input, then a small one with other sizes (stale keys / histograms of the first call lie in the array),
    # then the first again
    cases = [(70000, 400000, 5), (1000, 3000, 3), (70000, 400000, 5), (MAX_N, MAX_NZ, 8), (7, 30, 1)]
    grew = []
    for n, nz, seed in cases:
        ti, tj, tx = _triplets(n, nz, seed)
        ref = _run(L.spchol_triplet_to_csc_device, dev, n, ti, tj, tx)
        got = _run(L.spchol_triplet_to_csc_device_noalloc, dev, n, ti, tj, tx)
        assert ref[0][-1] == len(ref[1]) > 0
        assert np.array_equal(got[0], ref[0]) and np.array_equal(got[1], ref[1])
        assert np.array_equal(got[2], ref[2])                 # same additions in the same order: bit for bit
        grew.append(got[3])
    # device memory in use does not grow across a call once the kernels are loaded (first call: lazy module loading)
    assert all(g <= 0 for g in grew[1:]), grew


def test_noalloc_ingest_empty_and_bad_index():
    import torch
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    L = lib()
    z = np.zeros(0, np.int32)
    Ap, Ai, Ax, _ = _run(L.spchol_triplet_to_csc_device_noalloc, dev, 5, z, z, np.zeros(0))
    assert (Ap == 0).all() and len(Ai) == 0
    with pytest.raises(Exception):
        _run(L.spchol_triplet_to_csc_device_noalloc, dev, 3, np.array([0, 5], np.int32), np.array([0, 1], np.int32), np.array([1.0, 2.0]))
    # and the array is usable afterwards
    ti, tj, tx = _triplets(100, 5000, 2)
    ref = _run(L.spchol_triplet_to_csc_device, dev, 100, ti, tj, tx)
    got = _run(L.spchol_triplet_to_csc_device_noalloc, dev, 100, ti, tj, tx)
    assert all(np.array_equal(a, b) for a, b in zip(got[:3], ref[:3]))


def test_noalloc_ingest_refuses_sizes_beyond_the_static_work_space():
    L = lib()
    nnz = C.c_int64(0)
    dummy = C.c_void_p(256)                  # never dereferenced: the size check comes first
    for nz, n in ((MAX_NZ + 1, 10), (10, MAX_N + 1)):
        rc = L.spchol_triplet_to_csc_device_noalloc(nz, dummy, dummy, dummy, n, dummy, dummy, dummy, C.byref(nnz), None)
        assert rc == -5                       # SPCHOL_ERANGE
        msg = L.spchol_last_error().decode()
        assert "exceed the static work space" in msg and str(MAX_NZ) in msg
