"""The C ABI from plain C: tests/c_abi/consumer.c is compiled with gcc -std=c11 against include/slc.h
and linked against the shared library (no Python, no torch, no C++ on its side), and every result it
records is compared with the oracle.  The CPU part checks that the header is valid C and that every
prototype the consumer uses resolves against the library; the GPU part runs it."""
import os
import struct
import subprocess

import numpy as np
import pytest

import oracle
from helpers import arbitrary_triples
from sparse_linear_b200 import _capi
from sparse_linear_b200 import generators as g

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "c_abi", "consumer.c")


def build_consumer(tmp_path):
    exe = str(tmp_path / "consumer")
    libdir = os.path.dirname(_capi.LIB_PATH)
    cmd = ["gcc", "-std=c11", "-O1", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"),
           SRC, "-o", exe, "-L", libdir, "-lsparselinear_b200", "-Wl,-rpath," + libdir]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_header_is_c11_and_consumer_links(tmp_path):
    _capi.lib()
    exe = build_consumer(tmp_path)
    r = subprocess.run([exe, "--version"], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.strip() == "100", (r.stdout, r.stderr)


def _w(f, a, dtype):
    a = np.ascontiguousarray(a, dtype=dtype).ravel(order="K")
    f.write(struct.pack("<q", a.size))
    f.write(a.tobytes())


class Reader:
    def __init__(self, path):
        self.b = open(path, "rb").read()
        self.at = 0

    def arr(self, dtype):
        n = struct.unpack_from("<q", self.b, self.at)[0]
        self.at += 8
        a = np.frombuffer(self.b, dtype=dtype, count=n, offset=self.at).copy()
        self.at += n * np.dtype(dtype).itemsize
        return a

    def i64(self):
        return int(self.arr(np.int64)[0])

    def text(self):
        return self.arr(np.uint8).tobytes().decode()

    def matrix(self, dtype):
        nr, nc = self.i64(), self.i64()
        cp, ri = self.arr(np.int32), self.arr(np.int32)
        v = self.arr(np.float64)
        if np.dtype(dtype).kind == "c":
            v = v.view(np.complex128)
        return oracle.csc(nr, nc, cp, ri, v)


def same(a, b):
    return (a.nrows == b.nrows and a.ncols == b.ncols and np.array_equal(a.colptr, b.colptr)
            and np.array_equal(a.rowidx, b.rowidx) and np.array_equal(a.values, b.values))


def case(dtype, seed):
    rng = np.random.default_rng(seed)
    m = g.stencil3d_27pt((5, 6, 7), dtype)       # square, sliced-ELL plan, long rows (panel SpMM path)
    if np.dtype(dtype).kind == "c":
        m = g.CSC(m.nrows, m.ncols, m.colptr, m.rowidx, m.values * (1 + 0.5j) + 1j * rng.uniform(-1, 1, len(m.values)))
    k = 5
    ldx, ldy = m.ncols + 3, m.nrows + 2

    def rnd(*shape):
        a = rng.uniform(-1, 1, shape)
        if np.dtype(dtype).kind == "c":
            a = a + 1j * rng.uniform(-1, 1, shape)
        return a.astype(dtype)
    x, y0 = rnd(m.ncols), rnd(m.nrows)
    X = np.zeros((ldx, k), dtype=dtype, order="F")
    X[:m.ncols] = rnd(m.ncols, k)
    Y0 = np.zeros((ldy, k), dtype=dtype, order="F")
    Y0[:m.nrows] = rnd(m.nrows, k)
    tr, tc, tv = arbitrary_triples(rng, m.nrows, m.ncols, dtype, integer=False)
    return m, k, ldx, ldy, x, y0, X, Y0, tr[:4000], tc[:4000], tv[:4000]


@pytest.mark.gpu
def test_consumer_results_match_oracle(tmp_path):
    exe = build_consumer(tmp_path)
    cases = [case(np.float64, 1), case(np.complex128, 2)]
    fin, fout = str(tmp_path / "in.bin"), str(tmp_path / "out.bin")
    with open(fin, "wb") as f:
        for m, k, ldx, ldy, x, y0, X, Y0, tr, tc, tv in cases:
            _w(f, [m.nrows, m.ncols, k, ldx, ldy], np.int32)
            _w(f, m.colptr, np.int32)
            _w(f, m.rowidx, np.int32)
            _w(f, m.values.view(np.float64), np.float64)
            _w(f, x.view(np.float64), np.float64)
            _w(f, y0.view(np.float64), np.float64)
            _w(f, X.T.copy().view(np.float64), np.float64)      # column-major bytes
            _w(f, Y0.T.copy().view(np.float64), np.float64)
            _w(f, tr, np.int32)
            _w(f, tc, np.int32)
            _w(f, np.ascontiguousarray(tv).view(np.float64), np.float64)
    r = subprocess.run([exe, fin, fout], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    rd = Reader(fout)
    for m, k, ldx, ldy, x, y0, X, Y0, tr, tc, tv in cases:
        dt = m.values.dtype
        cx = dt.kind == "c"
        om = oracle.csc(*m)

        def vec():
            a = rd.arr(np.float64)
            return a.view(np.complex128) if cx else a

        def block():
            return vec().reshape((k, ldy)).T        # (ldy, k) column-major as written
        assert same(rd.matrix(dt), om)                                       # upload -> download
        want_mulv, want_axpy = oracle.mulv(om, x), oracle.axpy(om, x, y0)
        np.testing.assert_array_equal(vec(), want_mulv)                      # mulV, host
        np.testing.assert_array_equal(vec(), want_axpy)                      # axpy_, host
        np.testing.assert_array_equal(vec(), want_mulv)                      # device-resident
        np.testing.assert_array_equal(vec(), want_axpy)
        Xa = np.asfortranarray(X[:m.ncols])
        full = oracle.mulm(om, Xa)
        got = block()
        np.testing.assert_array_equal(got[:m.nrows], full)                   # mulM
        assert not got[m.nrows:].any()                                       # padding rows untouched
        got = block()                                                         # column range by pointer offset
        np.testing.assert_array_equal(got[:m.nrows, 1:k - 1], full[:, 1:k - 1])
        assert not got[:, 0].any() and not got[:, k - 1].any()
        acc = np.stack([oracle.axpy(om, np.ascontiguousarray(Xa[:, j]), np.ascontiguousarray(Y0[:m.nrows, j]))
                        for j in range(k)], axis=1)
        got = block()
        np.testing.assert_array_equal(got[:m.nrows], acc)                    # beta = 1: axpy_ per column, bit for bit
        np.testing.assert_array_equal(got[m.nrows:], Y0[m.nrows:])
        beta = (-0.5 + 0.25j) if cx else -0.5
        want_beta = np.stack([oracle.axpy(om, np.ascontiguousarray(Xa[:, j]), _scaled(beta, Y0[:m.nrows, j], cx))
                              for j in range(k)], axis=1)
        np.testing.assert_array_equal(block()[:m.nrows], want_beta)          # general beta: starts from beta * y
        np.testing.assert_array_equal(block()[:m.nrows], acc)                # device block, beta = 1
        assert rd.i64() == _capi.ERR_DIM
        assert rd.text().startswith("axpy_: column dimension")
        ot = oracle.transpose(om)
        oct_ = oracle.ctrans(om)
        assert same(rd.matrix(dt), ot)
        assert same(rd.matrix(dt), oct_)
        assert rd.i64() == int(oracle.hermitian(om))
        assert same(rd.matrix(dt), oracle.mm(ot, om))
        assert same(rd.matrix(dt), oracle.lin(1, om, 2, om))
        assert same(rd.matrix(dt), oracle.kronecker(ot, oct_))
        oa = oracle.compress(m.nrows, m.ncols, tr, tc, tv)
        assert same(rd.matrix(dt), oa)
        assert same(rd.matrix(dt), oa)
        assert rd.i64() == _capi.ERR_ROW_BOUNDS
        assert rd.i64() == 3
        assert "compress" in rd.text()
        np.testing.assert_array_equal(vec(), want_mulv)                      # slc_*_mulv_sharded (1 or 2 devices)
        rc = rd.i64()
        got = block()
        assert rc == 0
        np.testing.assert_array_equal(got[:m.nrows], full)                   # slc_*_mulm_sharded
    assert rd.i64() > 0                                                      # kernels of this library ran


def _scaled(beta, y, cx):
    """beta * y as the kernel forms it: one rounded product per entry (4-multiply formula for Complex)."""
    y = np.ascontiguousarray(y)
    if not cx:
        return beta * y
    br, bi = beta.real, beta.imag
    return (br * y.real - bi * y.imag) + 1j * (br * y.imag + bi * y.real)
