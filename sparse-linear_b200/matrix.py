"""Host-side mirror of the reference's Data.Matrix.Sparse interface for the hot path.

The reference is Haskell and no GHC exists in this image, so the host layer that tests and
benchmarks drive is this Python mirror over the C ABI (include/slc.h); the Haskell shim a
maintainer would add is in sparse-linear_b200/hs/ and INTEGRATION.md.  Names, argument
meaning and error messages follow /root/reference/sparse-linear/src/Data/Matrix/Sparse.hs:

    Matrix(ncols, nrows, pointers, indices, values)   Sparse.hs:67-76   (CSC)
    nonZero, cmap, scale                              Sparse.hs:115-125
    compress, decompress, fromTriples                 Sparse.hs:184-299,357-369
    transpose, ctrans, hermitian                      Sparse.hs:301-329,371-379
    axpy_, axpy, mulV, mulM                           Sparse.hs:433-498
    lin, +, -, *, mm                                  Sparse.hs:100-113,401-431,691-702
    kronecker, diag, ident, zeros                     Sparse.hs:597-679

Every product, transpose and assembly runs on the GPU through libsparselinear_b200.so.  There
is no CPU fallback: a missing library or device raises.
"""
import ctypes as C

import numpy as np

from . import _capi


class SparseLinearError(Exception):
    """The reference's `error` / `errorWithStackTrace` calls (e.g. Sparse.hs:252,453,702)."""

    def __init__(self, msg, code=None, pos=None):
        super().__init__(msg)
        self.code = code
        self.pos = pos


def _check(rc):
    if rc != _capi.OK:
        raise SparseLinearError(_capi.last_error(), code=rc)


def _prefix(dtype):
    return "z" if np.dtype(dtype).kind == "c" else "d"


def _elem_dtype(p):
    return np.complex128 if p == "z" else np.float64


def _fn(p, name):
    return getattr(_capi.lib(), "slc_%s_%s" % (p, name))


def _ptr(a):
    return C.c_void_p(a.ctypes.data)


class DeviceMatrix:
    """Owner of one slc_matrix handle (a ForeignPtr with finalizer &slc_free on the Haskell side)."""

    def __init__(self, handle):
        self.handle = handle

    def __del__(self):
        h, self.handle = getattr(self, "handle", None), None
        if h:
            try:
                _capi.lib().slc_free(h)
            except Exception:  # interpreter shutdown
                pass

    def dims(self):
        nr, nc, nz = C.c_int32(), C.c_int32(), C.c_int64()
        _check(_capi.lib().slc_dims(self.handle, C.byref(nr), C.byref(nc), C.byref(nz)))
        return nr.value, nc.value, nz.value

    def elem(self):
        return chr(_capi.lib().slc_elem_type(self.handle))

    def prepare(self):
        _check(_capi.lib().slc_prepare(self.handle))

    def set_spmv_kernel(self, kind):
        _check(_capi.lib().slc_set_spmv_kernel(self.handle, int(kind)))

    def plan_info(self):
        buf = C.create_string_buffer(256)
        _check(_capi.lib().slc_plan_info(self.handle, buf, 256))
        return buf.value.decode()


class Matrix:
    """CSC matrix: the five fields of the reference's record, plus a device copy kept beside it.

    A matrix that came out of a device operation downloads its arrays on first access.
    """

    def __init__(self, ncols, nrows, pointers, indices, values):
        self.ncols = int(ncols)
        self.nrows = int(nrows)
        values = np.ascontiguousarray(values)
        self._p = _prefix(values.dtype)
        self._pointers = np.ascontiguousarray(pointers, dtype=np.int32)
        self._indices = np.ascontiguousarray(indices, dtype=np.int32)
        self._values = np.ascontiguousarray(values, dtype=_elem_dtype(self._p))
        self._dev = None

    # -- construction from a device result ----------------------------------------------------
    @classmethod
    def _from_device(cls, handle):
        self = cls.__new__(cls)
        self._dev = DeviceMatrix(handle)
        self.nrows, self.ncols, self._nnz = self._dev.dims()
        self._p = self._dev.elem()
        self._pointers = self._indices = self._values = None
        return self

    @classmethod
    def _from_devmatrix(cls, dm):
        """A second view of an existing device matrix (shares the handle's owner)."""
        self = cls.__new__(cls)
        self._dev = dm
        self.nrows, self.ncols, self._nnz = dm.dims()
        self._p = dm.elem()
        self._pointers = self._indices = self._values = None
        return self

    def _download(self):
        if self._pointers is None:
            nnz = self._nnz
            cp = np.empty(self.ncols + 1, dtype=np.int32)
            ri = np.empty(nnz, dtype=np.int32)
            v = np.empty(nnz, dtype=_elem_dtype(self._p))
            _check(_fn(self._p, "download")(self._dev.handle, _ptr(cp), _ptr(ri), _ptr(v)))
            self._pointers, self._indices, self._values = cp, ri, v

    @property
    def pointers(self):
        self._download()
        return self._pointers

    @property
    def indices(self):
        self._download()
        return self._indices

    @property
    def values(self):
        self._download()
        return self._values

    @property
    def dtype(self):
        return np.dtype(_elem_dtype(self._p))

    def device(self):
        """Upload once (withConstMatrix's argument shape, Foreign.hs:24-41) and keep the handle."""
        if self._dev is None:
            h = C.c_void_p()
            _check(_fn(self._p, "upload")(self.nrows, self.ncols, _ptr(self._pointers), _ptr(self._indices),
                                          _ptr(self._values), C.byref(h)))
            self._dev = DeviceMatrix(h)
        return self._dev

    # -- derived Eq on every field (Sparse.hs:78) ------------------------------------------------
    def __eq__(self, other):
        if not isinstance(other, Matrix):
            return NotImplemented
        return (self.ncols == other.ncols and self.nrows == other.nrows and self._p == other._p
                and np.array_equal(self.pointers, other.pointers)
                and np.array_equal(self.indices, other.indices)
                and bool(np.all(self.values == other.values)))

    __hash__ = None

    def __repr__(self):
        return "Matrix {ncols = %d, nrows = %d, nonZero = %d, %s}" % (
            self.ncols, self.nrows, nonZero(self), self.dtype.name)

    # -- Num instance (Sparse.hs:100-113) ---------------------------------------------------------
    def __add__(self, other):
        return lin(1, self, 1, other)

    def __sub__(self, other):
        return lin(1, self, -1, other)

    def __mul__(self, other):
        return mm(self, other)

    def __neg__(self):
        return cmap(lambda v: -v, self)


# ---- accessors --------------------------------------------------------------------------------

def nonZero(m):
    """Sparse.hs:115-117"""
    if m._pointers is None:
        return m._nnz
    return int(m._pointers[-1])


def cmap(f, m):
    """Sparse.hs:119-121: map over the stored values, structure unchanged (host side)."""
    return Matrix(m.ncols, m.nrows, m.pointers, m.indices, f(m.values))


def scale(x, m):
    """Sparse.hs:123-125"""
    return cmap(lambda v: v * x, m)


def decompress(pointers):
    """Sparse.hs:293-299: one column id per stored entry."""
    pointers = np.asarray(pointers)
    return np.repeat(np.arange(len(pointers) - 1, dtype=np.int32), np.diff(pointers)).astype(np.int32)


def diag(values):
    """Sparse.hs:652-659"""
    values = np.ascontiguousarray(values)
    n = len(values)
    return Matrix(n, n, np.arange(n + 1), np.arange(n), values)


def ident(n, dtype=np.float64):
    """Sparse.hs:669-671"""
    return diag(np.ones(n, dtype=dtype))


def zeros(nrows, ncols, dtype=np.float64):
    """Sparse.hs:673-679"""
    return Matrix(ncols, nrows, np.zeros(ncols + 1, dtype=np.int32), np.zeros(0, np.int32),
                  np.zeros(0, dtype=dtype))


# ---- assembly ---------------------------------------------------------------------------------

def compress(nrows, ncols, rows, cols, vals):
    """COO -> CSC on the GPU (Sparse.hs:184-255)."""
    vals = np.ascontiguousarray(vals)
    if len(rows) != len(cols):
        raise SparseLinearError("compress: row and column array lengths differ")
    if len(rows) != len(vals):
        raise SparseLinearError("compress: row and value array lengths differ")
    p = _prefix(vals.dtype)

    def as_i32(a):
        # 32-bit index arrays (what the C ABI takes, Foreign.hs:24-41) are lent as they are; anything wider
        # is narrowed here, and indices beyond int32 -- which cannot be valid for a 32-bit matrix -- are
        # clamped so that they fail the bounds check
        if isinstance(a, np.ndarray) and a.dtype == np.int32:
            return np.ascontiguousarray(a)
        a = np.ascontiguousarray(a, dtype=np.int64)
        return np.clip(a, -1, np.iinfo(np.int32).max).astype(np.int32)
    r32, c32 = as_i32(rows), as_i32(cols)
    v = np.ascontiguousarray(vals, dtype=_elem_dtype(p))
    h = C.c_void_p()
    bad = C.c_int64(-1)
    rc = _fn(p, "compress")(int(nrows), int(ncols), len(r32), _ptr(r32), _ptr(c32), _ptr(v), C.byref(h),
                            C.byref(bad))
    if rc != _capi.OK:
        raise SparseLinearError(_capi.last_error(), code=rc, pos=bad.value if bad.value >= 0 else None)
    return Matrix._from_device(h)


def compress_dev(nrows, ncols, n, rows_ptr, cols_ptr, vals_ptr, dtype=np.float64):
    """compress over triplet arrays that already live in device memory (int32 rows / cols, values of
    `dtype`), given as raw device addresses; the arrays are read in place and left intact."""
    p = _prefix(np.dtype(dtype))
    h = C.c_void_p()
    bad = C.c_int64(-1)
    rc = _fn(p, "compress_dev")(int(nrows), int(ncols), int(n), C.c_void_p(rows_ptr), C.c_void_p(cols_ptr),
                                C.c_void_p(vals_ptr), C.byref(h), C.byref(bad))
    if rc != _capi.OK:
        raise SparseLinearError(_capi.last_error(), code=rc, pos=bad.value if bad.value >= 0 else None)
    return Matrix._from_device(h)


def fromTriples(nr, nc, triples, dtype=None):
    """Sparse.hs:357-363; triples are (row, col, value)."""
    triples = list(triples)
    rows = np.array([t[0] for t in triples], dtype=np.int64)
    cols = np.array([t[1] for t in triples], dtype=np.int64)
    vals = np.array([t[2] for t in triples])
    if dtype is not None:
        vals = vals.astype(dtype)
    elif vals.dtype.kind != "c":
        vals = vals.astype(np.float64)
    return compress(nr, nc, rows, cols, vals)


# ---- transpose ---------------------------------------------------------------------------------

def transpose(m):
    """Sparse.hs:301-329"""
    h = C.c_void_p()
    _check(_fn(m._p, "transpose")(m.device().handle, 0, C.byref(h)))
    return Matrix._from_device(h)


def ctrans(m):
    """Sparse.hs:371-375"""
    h = C.c_void_p()
    _check(_fn(m._p, "transpose")(m.device().handle, 1, C.byref(h)))
    return Matrix._from_device(h)


def hermitian(m):
    """Sparse.hs:377-379"""
    res = C.c_int(0)
    _check(_fn(m._p, "hermitian")(m.device().handle, C.byref(res)))
    return bool(res.value)


# ---- products ----------------------------------------------------------------------------------

def _vec(m, x, n, what):
    x = np.ascontiguousarray(x, dtype=m.dtype)
    if x.ndim != 1:
        raise SparseLinearError("axpy_: operand is not a vector")
    if what == "x" and len(x) != n:
        raise SparseLinearError("axpy_: column dimension %d does not match operand dimension %d" % (n, len(x)))
    if what == "y" and len(x) != n:
        raise SparseLinearError("axpy_: row dimension %d does not match result dimension %d" % (n, len(x)))
    return x


def axpy_(m, x, y):
    """y <- A x + y in place (Sparse.hs:433-453).  y must be a writable contiguous numpy vector."""
    xs = _vec(m, x, m.ncols, "x")
    if not (isinstance(y, np.ndarray) and y.dtype == m.dtype and y.flags.c_contiguous and y.flags.writeable):
        raise SparseLinearError("axpy_: result must be a writable contiguous %s vector" % m.dtype.name)
    _vec(m, y, m.nrows, "y")
    _check(_fn(m._p, "axpy")(m.device().handle, _ptr(xs), _ptr(y)))


def axpy(m, x, y):
    """A x + y (Sparse.hs:455-462)"""
    out = _vec(m, y, m.nrows, "y").copy()
    axpy_(m, x, out)
    return out


def mulV(m, x):
    """A x (Sparse.hs:464-471)"""
    xs = _vec(m, x, m.ncols, "x")
    y = np.empty(m.nrows, dtype=m.dtype)
    _check(_fn(m._p, "mulv")(m.device().handle, _ptr(xs), _ptr(y)))
    return y


def mulM(m, X, Y0=None, beta=None):
    """A X for a dense (ncols, k) block (mulM, Sparse.hs:473-498; FEAST's multiplyWork,
    feast/src/Numeric/LinearAlgebra/Feast.hs:203-208).  Returns an (nrows, k) Fortran-order array.
    With Y0 (and beta, default 1) the result is A X + beta Y0: beta = 1 is the reference's axpy_ into a
    block the caller filled (Sparse.hs:485-488)."""
    X = np.asarray(X)
    if X.ndim != 2 or X.shape[0] != m.ncols:
        raise SparseLinearError("mulM: inner dimension mismatch")
    Xf = np.asfortranarray(X, dtype=m.dtype)
    k = Xf.shape[1]
    bptr = None
    if Y0 is None:
        Y = np.empty((m.nrows, k), dtype=m.dtype, order="F")
    else:
        Y0 = np.asarray(Y0)
        if Y0.shape != (m.nrows, k):
            raise SparseLinearError("axpy_: row dimension %d does not match result dimension %d" % (m.nrows, Y0.shape[0]))
        Y = np.array(Y0, dtype=m.dtype, order="F", copy=True)
        b = np.array([1 if beta is None else beta], dtype=m.dtype)
        bptr = _ptr(b)
    _check(_fn(m._p, "mulm")(m.device().handle, _ptr(Xf), max(m.ncols, 1), _ptr(Y), max(m.nrows, 1), k, bptr))
    return Y


def mulV_dev(m, x_ptr, y_ptr, accumulate=False):
    """Device-resident product: x_ptr, y_ptr are raw device addresses; asynchronous."""
    _check(_fn(m._p, "axpy_dev" if accumulate else "mulv_dev")(m.device().handle, C.c_void_p(x_ptr),
                                                              C.c_void_p(y_ptr)))


def mulM_dev(m, X_ptr, ldx, Y_ptr, ldy, k, beta=None):
    bptr = None
    if beta is not None:
        b = np.array([beta], dtype=m.dtype)
        bptr = _ptr(b)
    _check(_fn(m._p, "mulm_dev")(m.device().handle, C.c_void_p(X_ptr), int(ldx), C.c_void_p(Y_ptr), int(ldy),
                                 int(k), bptr))


def _same(a, b, fn):
    if a._p != b._p:
        raise TypeError("%s: element types differ (%s vs %s)" % (fn, a.dtype.name, b.dtype.name))


def mm(a, b):
    """A B (Sparse.hs:691-702)"""
    _same(a, b, "mm")
    if a.ncols != b.nrows:
        raise SparseLinearError("mm: inner dimension mismatch")
    h = C.c_void_p()
    _check(_fn(a._p, "mm")(a.device().handle, b.device().handle, C.byref(h)))
    return Matrix._from_device(h)


def lin(alpha, a, beta, b):
    """alpha A + beta B on the union pattern (Sparse.hs:426-431)"""
    _same(a, b, "glin")
    if a.nrows != b.nrows:
        raise SparseLinearError("glin: row number mismatch")
    if a.ncols != b.ncols:
        raise SparseLinearError("glin: column number mismatch")
    al = np.array([alpha], dtype=a.dtype)
    be = np.array([beta], dtype=a.dtype)
    h = C.c_void_p()
    _check(_fn(a._p, "lin")(_ptr(al), a.device().handle, _ptr(be), b.device().handle, C.byref(h)))
    return Matrix._from_device(h)


def kronecker(a, b):
    """Sparse.hs:597-638"""
    _same(a, b, "kronecker")
    h = C.c_void_p()
    _check(_fn(a._p, "kronecker")(a.device().handle, b.device().handle, C.byref(h)))
    return Matrix._from_device(h)


# ---- library state ------------------------------------------------------------------------------

def set_device(dev):
    _check(_capi.lib().slc_set_device(int(dev)))


def set_stream(stream_ptr):
    _check(_capi.lib().slc_set_stream(C.c_void_p(stream_ptr)))


def synchronize():
    _check(_capi.lib().slc_synchronize())


def launch_count():
    return int(_capi.lib().slc_launch_count())


def spmm_window_error():
    """Nonzero if a launch of the TMA-staged block product gave up on a barrier (diagnostic)."""
    f = C.c_int(0)
    _check(_capi.lib().slc_spmm_window_error(C.byref(f)))
    return f.value


def device_count():
    n = C.c_int(0)
    rc = _capi.lib().slc_device_count(C.byref(n))
    return n.value if rc == _capi.OK else 0
