"""Matrix Market parse on the GPU (csrc/mtx_parse.cuh, spchol_mtx_read_gpu / spchol_mtx_parse_device) against the product's
host reader -- the restatement of the reference's load_matrix_market_to_csc (include/mtx_reader.hpp:17-62), itself checked
against the reference in tests/test_symbolic.py.  Bit for bit: same triplets, same CSC."""
import ctypes as C
import os

import numpy as np
import pytest

import sparsecholesky_b200 as sc
from sparsecholesky_b200 import synth
from sparsecholesky_b200._lib import check, lib, ptr

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def host_read(path):
    L = lib()
    n = C.c_int(0); nz = C.c_int64(0)
    check(L.spchol_mtx_read(path.encode(), C.byref(n), C.byref(nz), None, None, None), "mtx_read")
    ti = np.empty(nz.value, np.int32); tj = np.empty(nz.value, np.int32); tx = np.empty(nz.value)
    check(L.spchol_mtx_read(path.encode(), C.byref(n), C.byref(nz), ptr(ti), ptr(tj), ptr(tx)), "mtx_read")
    return n.value, ti, tj, tx


def gpu_read(path):
    L = lib()
    n = C.c_int(0); nz = C.c_int64(0); ms = C.c_double(0)
    check(L.spchol_mtx_read_gpu(path.encode(), C.byref(n), C.byref(nz), None, None, None, None), "mtx_read_gpu")
    ti = np.full(nz.value, -9, np.int32); tj = np.full(nz.value, -9, np.int32); tx = np.full(nz.value, np.nan)
    check(L.spchol_mtx_read_gpu(path.encode(), C.byref(n), C.byref(nz), ptr(ti), ptr(tj), ptr(tx), C.byref(ms)), "mtx_read_gpu")
    return n.value, ti, tj, tx, ms.value


def same_bits(a, b):
    return np.array_equal(np.asarray(a, np.float64).view(np.uint64), np.asarray(b, np.float64).view(np.uint64))


@pytest.mark.parametrize("name", ["bcsstk01", "1138_bus"])
def test_reference_matrices(name):
    path = os.path.join(ROOT, "tests", "data", name + ".mtx")
    n, hi, hj, hx = host_read(path)
    n2, gi, gj, gx, _ = gpu_read(path)
    assert n2 == n and np.array_equal(gi, hi) and np.array_equal(gj, hj) and same_bits(gx, hx)
    A = sc.load_matrix_market_to_csc(path)
    B = sc.load_matrix_market_to_csc_gpu(path)
    assert B.n == A.n and np.array_equal(B.p, A.p) and np.array_equal(B.i, A.i) and same_bits(B.x, A.x)


def test_free_form_text_and_values_for_the_host(tmp_path):
    rng = np.random.default_rng(5)
    nz, n = 4000, 977
    r = rng.integers(1, n + 1, nz); c = rng.integers(1, n + 1, nz)
    v = rng.standard_normal(nz) * 10.0 ** rng.integers(-30, 30, nz)
    seps = [" ", "\n", "\t", "\r\n", "  ", " \t ", "\n\n", "\f", "\v "]
    parts = [f"{n}", f"+{n}", f"{nz}"]
    for k in range(nz):
        val = "%.17g" % v[k] if k % 3 else "%+.16E" % v[k]
        if k == 17:
            val = "1.000000000000000000000001e5"            # 25 digits: the host's strtod converts it
        if k == 18:
            val = "-inf"
        parts += [str(r[k]) if k % 5 else "+" + str(r[k]), str(c[k]), val]
    text = "".join(p + seps[rng.integers(0, len(seps))] for p in parts) + "  trailing tokens are never read 1 2 3"
    f = tmp_path / "free.mtx"
    f.write_bytes(b"%%MatrixMarket matrix coordinate real symmetric\n% a comment\n" + text.encode())
    n1, hi, hj, hx = host_read(str(f))
    n2, gi, gj, gx, _ = gpu_read(str(f))
    assert n1 == n2 == n and np.array_equal(gi, hi) and np.array_equal(gj, hj) and same_bits(gx, hx)
    assert gx[17] == 100000.0 and gx[18] == -np.inf


def test_large_file_text_to_csc_without_leaving_the_device(tmp_path):
    """512^2 Laplacian with random values as a 25 MB coordinate file: spchol_mtx_parse_device -> spchol_triplet_to_csc_device
    on device buffers (torch: device memory only), compared with the host reader + host conversion."""
    import torch
    dims = (512, 512, 1)
    ti, tj, tx, n = synth.laplacian_triplets(*dims, perm=synth.nd_perm(*dims))
    rng = np.random.default_rng(3)
    tx = tx * (1.0 + rng.random(len(tx)))
    swap = rng.random(len(ti)) < 0.5                        # half of the entries written as (col, row): the reader mirrors them
    r = np.where(swap, tj, ti) + 1; c = np.where(swap, ti, tj) + 1
    f = tmp_path / "lap.mtx"
    with open(f, "w") as fh:
        fh.write("%%MatrixMarket matrix coordinate real symmetric\n%\n")
        fh.write(f"{n} {n} {len(ti)}\n")
        fh.write("".join("%d %d %.17g\n" % t for t in zip(r.tolist(), c.tolist(), tx.tolist())))
    n1, hi, hj, hx = host_read(str(f))
    n2, gi, gj, gx, ms = gpu_read(str(f))
    print(f"device Matrix Market parse: {os.path.getsize(f) / 1e6:.1f} MB, {len(hi)} entries in {ms:.2f} ms")
    assert n1 == n2 == n and np.array_equal(gi, hi) and np.array_equal(gj, hj) and same_bits(gx, hx) and same_bits(gx, tx)
    # the same on device buffers, straight into the device COO -> CSC
    raw = open(f, "rb").read()
    k = 0
    while raw[k:k + 1] == b"%":
        k = raw.index(b"\n", k) + 1
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    d_text = torch.frombuffer(bytearray(raw[k:]), dtype=torch.uint8).to(dev)
    nz = len(hi)
    d_ti = torch.empty(nz, dtype=torch.int32, device=dev); d_tj = torch.empty(nz, dtype=torch.int32, device=dev)
    d_tx = torch.empty(nz, dtype=torch.float64, device=dev)
    L = lib()
    nn = C.c_int(0); nnz = C.c_int64(0); hc = C.c_int64(-1)
    check(L.spchol_mtx_parse_device(d_text.data_ptr(), d_text.numel(), nz, d_ti.data_ptr(), d_tj.data_ptr(), d_tx.data_ptr(),
                                    C.byref(nn), C.byref(nnz), C.byref(hc), None), "mtx_parse_device")
    assert nn.value == n and nnz.value == nz and hc.value == 0
    d_Ap = torch.empty(n + 1, dtype=torch.int32, device=dev); d_Ai = torch.empty(nz, dtype=torch.int32, device=dev)
    d_Ax = torch.empty(nz, dtype=torch.float64, device=dev)
    m = C.c_int64(0)
    check(L.spchol_triplet_to_csc_device(nz, d_ti.data_ptr(), d_tj.data_ptr(), d_tx.data_ptr(), n, d_Ap.data_ptr(), d_Ai.data_ptr(),
                                         d_Ax.data_ptr(), C.byref(m), None), "triplet_to_csc_device")
    A = sc.triplet_to_csc_matrix(hi, hj, hx, n)
    assert m.value == len(A.i) and np.array_equal(d_Ap.cpu().numpy(), A.p) and np.array_equal(d_Ai[:m.value].cpu().numpy(), A.i)
    assert same_bits(d_Ax[:m.value].cpu().numpy(), A.x)


def test_errors(tmp_path):
    def write(name, body):
        p = tmp_path / name
        p.write_bytes(b"%%MatrixMarket matrix coordinate real symmetric\n" + body)
        return str(p)
    L = lib()
    for body in (b"3 3 2\n1 1 4.0\n2 1\n", b"3 3 2\n1 x 4.0\n2 1 1.0\n", b"3 3 2\n1 1 4.0.1\n2 1 1.0\n", b"3 3 2\n1 1 abc\n2 1 1.0\n"):
        path = write("bad.mtx", body)
        n = C.c_int(0); nz = C.c_int64(0)
        ti = np.zeros(2, np.int32); tj = np.zeros(2, np.int32); tx = np.zeros(2)
        assert L.spchol_mtx_read(path.encode(), C.byref(n), C.byref(nz), ptr(ti), ptr(tj), ptr(tx)) < 0
        assert L.spchol_mtx_read_gpu(path.encode(), C.byref(n), C.byref(nz), ptr(ti), ptr(tj), ptr(tx), None) < 0
    with pytest.raises(RuntimeError):
        sc.load_matrix_market_to_csc_gpu("/nonexistent/file.mtx")
    ok = write("ok.mtx", b"3 3 2\n1 1 4.0\n3 1 -1.0\n")
    A = sc.load_matrix_market_to_csc_gpu(ok)
    assert A.n == 3 and list(A.p) == [0, 1, 1, 2] and list(A.i) == [0, 0] and list(A.x) == [4.0, -1.0]
