"""The device-side Matrix Market parse (csrc/mtx_parse.cuh) on the host: tests/cpp/mtx_emu.cpp runs the kernels' source chunk
after chunk.  Decimal -> double against Python's float() (correctly rounded), the token stream against the product's host
reader (spchol_mtx_read, the restatement of the reference's `ifstream >>` loop, include/mtx_reader.hpp:17-62)."""
import ctypes as C
import os
import random
import subprocess

import numpy as np
import pytest

import sparsecholesky_b200 as sc
from sparsecholesky_b200._lib import lib, ptr

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "sparsecholesky_b200", "csrc")
TOK_MAX = 64


@pytest.fixture(scope="module")
def emu():
    so = os.path.join(ROOT, "tests", "cpp", "libmtxemu.so")
    src = os.path.join(ROOT, "tests", "cpp", "mtx_emu.cpp")
    deps = [src, os.path.join(CSRC, "mtx_parse.cuh"), os.path.join(CSRC, "mtx_pow5.inc")]
    if not os.path.exists(so) or any(os.path.getmtime(d) > os.path.getmtime(so) for d in deps):
        subprocess.run(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-I", CSRC, src, "-o", so], check=True)
    L = C.CDLL(so)
    L.mtxemu_reals.argtypes = [C.c_char_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
    L.mtxemu_parse.restype = C.c_int64
    L.mtxemu_parse.argtypes = [C.c_char_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                               C.c_void_p, C.c_uint, C.c_void_p, C.c_void_p, C.c_int]
    return L


def reals(L, tokens):
    text = ("\n".join(tokens) + "\n").encode()
    off = np.zeros(len(tokens) + 1, np.int64)
    off[1:] = np.cumsum(np.fromiter((len(t) + 1 for t in tokens), np.int64, len(tokens)))
    st = np.zeros(len(tokens), np.int32); val = np.zeros(len(tokens))
    L.mtxemu_reals(text, off.ctypes.data, len(tokens), st.ctypes.data, val.ctypes.data)
    return st, val


def same_bits(a, b):
    return np.array_equal(np.asarray(a, np.float64).view(np.uint64), np.asarray(b, np.float64).view(np.uint64))


def test_known_hard_decimals(emu):
    toks = ["9007199254740993", "9007199254740992", "9007199254740995", "1e23", "8.41e21", "2.2250738585072011e-308",
            "2.2250738585072014e-308", "4.9e-324", "2.4703282292062327e-324", "2.4703282292062328e-324", "1.7976931348623157e308",
            "1.7976931348623159e308", "1e309", "1e-400", "0.000000000000000000000000001", "-0.0", "0", "+1.5", "1.", "0.1", "1e0",
            "1234567890123456789", "6.0221409e+23", "1.0000000000000000e+00", "-2.8322685185200000e+06", "5e-324", "3e-324", "2e-324",
            "1e22", "1e-22", "7.2057594037927933e16", "17976931348623157e292", ".5", "5.e-1", "1E5", "1e+5", "1.7976931348623158e308",
            "4503599627370496.5", "4503599627370497.5", "0.3", "-1e-5", "00012.5000"]
    st, val = reals(emu, toks)
    assert (st == 0).all()
    assert same_bits(val, [float(t) for t in toks])
    assert np.signbit(val[toks.index("-0.0")])
    # left to the host (status 1) and refused (status 2)
    st, _ = reals(emu, ["12345678901234567890", "0.500000000000000166533453693773481063544750213623046875", "inf", "-nan", "NaN", "e5"])
    assert (st == 1).all()
    st, _ = reals(emu, ["1.0abc", ".", "-", "1e", "1e+", "1.0D+00", "--1", "1,5", "0x10", "1..2", "+"])
    assert (st == 2).all()


def test_decimal_to_double_is_correctly_rounded(emu):
    rng = np.random.default_rng(1)
    random.seed(2)
    toks = []
    d = rng.integers(0, 2 ** 63, 60000, dtype=np.uint64).view(np.float64)       # every exponent, subnormals included
    d = d[np.isfinite(d)]
    for fmt in ("%.17g", "%.16e", "%.15g", "%.6g", "%.18e"):
        toks += [fmt % x for x in d[:30000]]
    d2 = rng.standard_normal(40000) * 10.0 ** rng.integers(-12, 12, 40000)
    toks += ["%.17g" % x for x in d2] + ["%.16e" % x for x in d2] + [repr(float(x)) for x in d2]
    for _ in range(80000):                                                       # any 1..19 digits, any exponent
        nd = random.randint(1, 19); w = random.randint(10 ** (nd - 1), 10 ** nd - 1)
        toks.append(f"{w}e{random.randint(-345, 310)}")
        if nd > 1:
            k = random.randint(1, nd - 1); s = str(w)
            toks.append(f"{s[:k]}.{s[k:]}E{random.randint(-330, 300):+d}")
    for _ in range(40000):                                                       # exact half-way cases and their neighbours
        M = random.getrandbits(52) | (1 << 52)
        v = (2 * M + 1) << (random.randint(1, 9) - 1)
        for dv in (-1, 0, 1):
            if v + dv < 10 ** 19:
                toks.append(str(v + dv)); toks.append(str(v + dv) + "e" + str(random.randint(-3, 3)))
        k = random.randint(1, 4); num = (2 * M + 1) * 5 ** k
        if num < 10 ** 19:
            s = str(num); toks.append(s[:-k] + "." + s[-k:]); toks.append(s + "e-" + str(k))
    toks += [f"1e{q}" for q in range(-350, 320)] + [f"9.999999999999999999e{q}" for q in range(-330, 309)]
    st, val = reals(emu, toks)
    assert (st == 0).all()                      # (the "cannot tell" case of the conversion is so rare that none shows up here)
    assert same_bits(val, [float(t) for t in toks])


def parse(L, text, nz_cap, mirror=1, fb_cap=1024):
    ti = np.full(max(nz_cap, 1), -9, np.int32); tj = np.full(max(nz_cap, 1), -9, np.int32); tx = np.full(max(nz_cap, 1), np.nan)
    hdr = np.zeros(3, np.int64); err = np.zeros(1, np.uint64); ntok = np.zeros(1, np.uint64); fbn = np.zeros(1, np.uint32)
    fbe = np.zeros(fb_cap, np.int32); fbt = np.zeros(fb_cap * TOK_MAX, np.uint8)
    L.mtxemu_parse(text, len(text), nz_cap, ti.ctypes.data, tj.ctypes.data, tx.ctypes.data, hdr.ctypes.data, err.ctypes.data,
                   ntok.ctypes.data, fbn.ctypes.data, fb_cap, fbe.ctypes.data, fbt.ctypes.data, mirror)
    fb = {}
    for k in range(min(int(fbn[0]), fb_cap)):
        raw = bytes(fbt[k * TOK_MAX:(k + 1) * TOK_MAX])
        fb[int(fbe[k])] = raw[:raw.index(0)].decode()
    return ti, tj, tx, hdr, int(err[0]), int(ntok[0]), fb


def host_read(path):
    Lh = lib()
    n = C.c_int(0); nz = C.c_int64(0)
    assert Lh.spchol_mtx_read(path.encode(), C.byref(n), C.byref(nz), None, None, None) == 0
    ti = np.empty(nz.value, np.int32); tj = np.empty(nz.value, np.int32); tx = np.empty(nz.value)
    assert Lh.spchol_mtx_read(path.encode(), C.byref(n), C.byref(nz), ptr(ti), ptr(tj), ptr(tx)) == 0
    return n.value, ti, tj, tx


def data_part(path):
    raw = open(path, "rb").read()
    k = 0
    while k < len(raw) and raw[k:k + 1] == b"%":
        k = raw.index(b"\n", k) + 1
    return raw[k:]


NOMORE = 2 ** 64 - 1


@pytest.mark.parametrize("name", ["bcsstk01", "1138_bus"])
def test_reference_matrices_token_stream(emu, name):
    path = os.path.join(ROOT, "tests", "data", name + ".mtx")
    n, hi, hj, hx = host_read(path)
    text = data_part(path)
    ti, tj, tx, hdr, err, ntok, fb = parse(emu, text, len(hi))
    assert err == NOMORE and not fb and list(hdr) == [n, n, len(hi)] and ntok == 3 + 3 * len(hi)
    assert np.array_equal(ti, hi) and np.array_equal(tj, hj) and same_bits(tx, hx)


def test_free_form_text_long_tokens_and_chunk_borders(emu, tmp_path):
    # tokens separated by any whitespace, spread over lines at will, CRLF, tabs, signs; values of 17 digits; one of 25 digits
    # and an "inf" for the host; entries made so that tokens straddle the 256-byte chunk borders at every phase
    rng = np.random.default_rng(5)
    nz, n = 4000, 977
    r = rng.integers(1, n + 1, nz); c = rng.integers(1, n + 1, nz)
    v = rng.standard_normal(nz) * 10.0 ** rng.integers(-30, 30, nz)
    seps = [" ", "\n", "\t", "\r\n", "  ", " \t ", "\n\n", "\f", "\v "]
    parts = [f"{n}", f"+{n}", f"{nz}"]
    for k in range(nz):
        val = "%.17g" % v[k] if k % 3 else "%+.16E" % v[k]
        if k == 17:
            val = "1.000000000000000000000001e5"
        if k == 18:
            val = "-inf"
        parts += [str(r[k]) if k % 5 else "+" + str(r[k]), str(c[k]), val]
    text = ""
    for ptxt in parts:
        text += ptxt + seps[rng.integers(0, len(seps))]
    text = text + "  trailing tokens are never read 1 2 3"
    f = tmp_path / "free.mtx"
    f.write_bytes(b"%%MatrixMarket matrix coordinate real symmetric\n% a comment\n" + text.encode())
    n2, hi, hj, hx = host_read(str(f))
    ti, tj, tx, hdr, err, ntok, fb = parse(emu, text.encode(), nz)
    assert n2 == n and list(hdr) == [n, n, nz] and ntok > 3 + 3 * nz
    assert err >= 3 + 3 * nz                                  # the words after the entries are not numbers, and nobody cares
    assert fb == {17: "1.000000000000000000000001e5", 18: "-inf"}
    for e, t in fb.items():
        tx[e] = float(t)
    assert np.array_equal(ti, hi) and np.array_equal(tj, hj) and same_bits(tx, hx)
    # a larger capacity than the file has entries: the surplus slots stay untouched apart from the trailing tokens' places
    ti2, tj2, tx2, hdr2, err2, _, _ = parse(emu, text.encode(), nz + 2)
    assert np.array_equal(ti2[:nz], hi) and err2 >= 3 + 3 * nz


def test_malformed_and_truncated(emu):
    ok = b"3 3 2\n1 1 4.0\n2 1 -1.0\n"
    ti, tj, tx, hdr, err, ntok, fb = parse(emu, ok, 2)
    assert err == NOMORE and ntok == 9 and list(ti) == [0, 0] and list(tj) == [0, 1] and list(tx) == [4.0, -1.0]   # mirrored
    ti, tj, *_ = parse(emu, ok, 2, mirror=0)
    assert list(ti) == [0, 1] and list(tj) == [0, 0]
    assert parse(emu, b"3 3 2\n1 1 4.0\n2 1\n", 2)[5] == 8                      # truncated: 8 tokens < 3 + 3 * 2
    assert parse(emu, b"3 3 2\n1 x 4.0\n2 1 1.0\n", 2)[4] == 4                  # token 4 is not an integer
    assert parse(emu, b"3 3 2\n1 1 4.0.1\n2 1 1.0\n", 2)[4] == 5
    assert parse(emu, b"3 3.0 2\n1 1 4.0\n2 1 1.0\n", 2)[4] == 1
    assert parse(emu, b"3 3 2\n1 1.5 4.0\n2 1 1.0\n", 2)[4] == 4
    assert parse(emu, b"3 3 2\n1 99999999999 4.0\n2 1 1.0\n", 2)[4] == 4        # beyond int
    big = b"2 2 1\n1 1 " + b"1" * 70 + b"\n"                                    # a value token too long to hand over
    assert parse(emu, big, 1)[4] == 5
