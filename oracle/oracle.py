"""ctypes binding of oracle/_build/liboracle.so (the C restatement of the reference).

TEST INFRASTRUCTURE ONLY -- see oracle/oracle.h.  Matrices are passed as
(nrows, ncols, colptr int32, rowidx int32, values float64|complex128), the convention of
/root/reference/sparse-linear/src/Data/Matrix/Sparse/Foreign.hs:24-41.
"""
import ctypes as C
import os
import subprocess
from collections import namedtuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "liboracle.so")

OK, ERR_DIM, ERR_ROW_BOUNDS, ERR_COL_BOUNDS, ERR_ALLOC = 0, -1, -2, -3, -4

CSC = namedtuple("CSC", "nrows ncols colptr rowidx values")


class OracleError(Exception):
    def __init__(self, code, msg, pos=None):
        super().__init__(msg)
        self.code = code
        self.pos = pos


class _OrcCsc(C.Structure):
    _fields_ = [("nrows", C.c_int32), ("ncols", C.c_int32), ("nnz", C.c_int64),
                ("colptr", C.POINTER(C.c_int32)), ("rowidx", C.POINTER(C.c_int32)),
                ("vals", C.POINTER(C.c_double))]


def build(force=False):
    """Compile the oracle (gcc, a second or two).  Building the checker is not using it."""
    srcs = [os.path.join(_HERE, f) for f in ("oracle.c", "oracle_impl.inc", "oracle.h", "Makefile")]
    if (not force and os.path.exists(_LIB_PATH)
            and os.path.getmtime(_LIB_PATH) >= max(os.path.getmtime(s) for s in srcs)):
        return _LIB_PATH
    subprocess.run(["make", "-C", _HERE, "-B", "CC=gcc"], check=True, capture_output=True)
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.orc_csc_free.argtypes = [C.POINTER(_OrcCsc)]
        _lib.orc_csc_free.restype = None
    return _lib


def _prefix(values):
    return "z" if np.iscomplexobj(values) else "d"


def _dtype(p):
    return np.complex128 if p == "z" else np.float64


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _vals(a, p):
    return np.ascontiguousarray(a, dtype=_dtype(p))


def _ptr(a, ty):
    return a.ctypes.data_as(C.POINTER(ty))


def _mat_args(m, p):
    cp, ri, v = _i32(m.colptr), _i32(m.rowidx), _vals(m.values, p)
    keep = (cp, ri, v)
    return [C.c_int32(m.nrows), C.c_int32(m.ncols), _ptr(cp, C.c_int32), _ptr(ri, C.c_int32),
            _ptr(v, C.c_double)], keep


def _take(out, p):
    nnz = int(out.nnz)
    cp = np.ctypeslib.as_array(out.colptr, shape=(out.ncols + 1,)).copy()
    ri = np.ctypeslib.as_array(out.rowidx, shape=(max(nnz, 1),))[:nnz].copy()
    w = 2 if p == "z" else 1
    raw = np.ctypeslib.as_array(out.vals, shape=(max(nnz * w, 1),))[:nnz * w].copy()
    v = raw.view(np.complex128) if p == "z" else raw
    res = CSC(int(out.nrows), int(out.ncols), cp, ri, v)
    lib().orc_csc_free(C.byref(out))
    return res


def _same_type(*mats):
    p = "z" if any(np.iscomplexobj(m.values) for m in mats) else "d"
    return p


def csc(nrows, ncols, colptr, rowidx, values):
    return CSC(int(nrows), int(ncols), _i32(colptr), _i32(rowidx), np.ascontiguousarray(values))


# ---- SpMV / SpMM ----------------------------------------------------------------------------

def mulv(m, x):
    p = _same_type(m)
    args, keep = _mat_args(m, p)
    x = _vals(x, p)
    y = np.empty(m.nrows, dtype=_dtype(p))
    getattr(lib(), "orc_%s_mulv" % p)(*args, _ptr(x, C.c_double), _ptr(y, C.c_double))
    return y


def axpy(m, x, y):
    p = _same_type(m)
    args, keep = _mat_args(m, p)
    x = _vals(x, p)
    y = _vals(y, p).copy()
    getattr(lib(), "orc_%s_axpy" % p)(*args, _ptr(x, C.c_double), _ptr(y, C.c_double))
    return y


def mulm(m, X):
    """X: (ncols, k) array; returns (nrows, k).  Column-major inside, like FEAST's work arrays."""
    p = _same_type(m)
    args, keep = _mat_args(m, p)
    Xf = np.asfortranarray(X, dtype=_dtype(p))
    k = Xf.shape[1]
    Y = np.empty((m.nrows, k), dtype=_dtype(p), order="F")
    getattr(lib(), "orc_%s_mulm" % p)(*args, _ptr(Xf, C.c_double), C.c_int64(max(m.ncols, 1)),
                                      _ptr(Y, C.c_double), C.c_int64(max(m.nrows, 1)), C.c_int32(k))
    return Y


# ---- transpose ------------------------------------------------------------------------------

def transpose(m, conj=False):
    p = _same_type(m)
    args, keep = _mat_args(m, p)
    out = _OrcCsc()
    rc = getattr(lib(), "orc_%s_transpose" % p)(*args, C.c_int(1 if conj else 0), C.byref(out))
    if rc:
        raise OracleError(rc, "transpose failed")
    return _take(out, p)


def ctrans(m):
    return transpose(m, conj=True)


def hermitian(m):
    p = _same_type(m)
    args, keep = _mat_args(m, p)
    return bool(getattr(lib(), "orc_%s_hermitian" % p)(*args))


# ---- SpGEMM / lin / kronecker ------------------------------------------------------------------

def mm(a, b, literal=False):
    p = _same_type(a, b)
    aa, k1 = _mat_args(a, p)
    bb, k2 = _mat_args(b, p)
    out = _OrcCsc()
    fn = getattr(lib(), "orc_%s_%s" % (p, "mm_literal" if literal else "mm"))
    rc = fn(*aa, *bb, C.byref(out))
    if rc == ERR_DIM:
        raise OracleError(rc, "mm: inner dimension mismatch")
    if rc:
        raise OracleError(rc, "mm failed")
    return _take(out, p)


def lin(alpha, a, beta, b, literal=False):
    p = _same_type(a, b)
    aa, k1 = _mat_args(a, p)
    bb, k2 = _mat_args(b, p)
    al = np.array([alpha], dtype=_dtype(p))
    be = np.array([beta], dtype=_dtype(p))
    out = _OrcCsc()
    fn = getattr(lib(), "orc_%s_%s" % (p, "lin_literal" if literal else "lin"))
    rc = fn(_ptr(al, C.c_double), *aa, _ptr(be, C.c_double), *bb, C.byref(out))
    if rc == ERR_DIM:
        which = "row" if a.nrows != b.nrows else "column"
        raise OracleError(rc, "glin: %s number mismatch" % which)
    if rc:
        raise OracleError(rc, "lin failed")
    return _take(out, p)


def add(a, b):
    return lin(1, a, 1, b)


def kronecker(a, b):
    p = _same_type(a, b)
    aa, k1 = _mat_args(a, p)
    bb, k2 = _mat_args(b, p)
    out = _OrcCsc()
    rc = getattr(lib(), "orc_%s_kronecker" % p)(*aa, *bb, C.byref(out))
    if rc:
        raise OracleError(rc, "kronecker failed")
    return _take(out, p)


# ---- assembly -----------------------------------------------------------------------------------

def compress(nrows, ncols, rows, cols, vals):
    vals = np.ascontiguousarray(vals)
    p = "z" if np.iscomplexobj(vals) else "d"
    rows, cols, vals = _i32(rows), _i32(cols), _vals(vals, p)
    if len(rows) != len(cols):
        raise OracleError(ERR_DIM, "compress: row and column array lengths differ")
    if len(rows) != len(vals):
        raise OracleError(ERR_DIM, "compress: row and value array lengths differ")
    out = _OrcCsc()
    bad = C.c_int64(-1)
    rc = getattr(lib(), "orc_%s_compress" % p)(C.c_int32(nrows), C.c_int32(ncols), C.c_int64(len(rows)),
                                               _ptr(rows, C.c_int32), _ptr(cols, C.c_int32),
                                               _ptr(vals, C.c_double), C.byref(out), C.byref(bad))
    if rc == ERR_ROW_BOUNDS:
        raise OracleError(rc, "compress: row index out of bounds (0,%d) at %d" % (nrows, bad.value), bad.value)
    if rc == ERR_COL_BOUNDS:
        raise OracleError(rc, "compress: column index out of bounds (0,%d) at %d" % (ncols, bad.value), bad.value)
    if rc:
        raise OracleError(rc, "compress failed")
    return _take(out, p)


def from_triples(nrows, ncols, triples):
    rows = [t[0] for t in triples]
    cols = [t[1] for t in triples]
    vals = [t[2] for t in triples]
    return compress(nrows, ncols, rows, cols, vals)


def decompress(colptr):
    colptr = _i32(colptr)
    out = np.empty(int(colptr[-1]), dtype=np.int32)
    lib().orc_decompress(C.c_int32(len(colptr) - 1), _ptr(colptr, C.c_int32), _ptr(out, C.c_int32))
    return out


def check_format(m):
    cp, ri = _i32(m.colptr), _i32(m.rowidx)
    if len(cp) != m.ncols + 1:
        return 2
    if len(ri) != len(m.values):
        return 4
    return int(lib().orc_check_format(C.c_int32(m.nrows), C.c_int32(m.ncols), C.c_int64(len(ri)),
                                      _ptr(cp, C.c_int32), _ptr(ri, C.c_int32)))


def ident(n, dtype=np.float64):
    return csc(n, n, np.arange(n + 1), np.arange(n), np.ones(n, dtype=dtype))


def diag(v):
    v = np.ascontiguousarray(v)
    n = len(v)
    return csc(n, n, np.arange(n + 1), np.arange(n), v)
