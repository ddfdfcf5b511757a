"""FEAST adapter stand-in: the data flow of the reference's reverse-communication loop around the one
call this library serves, `multiplyWork` (feast/src/Numeric/LinearAlgebra/Feast.hs:189-208).

FEAST itself (a Fortran library the reference binds, Feast/Internal.hs:24-43) is not in this image, so
what is reproduced is everything on the Haskell side of `feast_rci` that touches the hot path:

  * `_eigenvectors`, `_work1` are host vectors of m0*n elements, column-major, sliced into columns by
    `sliceMat` (Feast.hs:189-192);
  * on ijob 30 (matA) / 40 (matB) `multiplyWork mat` reads fpm[23], fpm[24] (1-based first column and
    count of the active range) and does, for each active column c,
    `MV.set work1[c] 0 >> axpy_ mat eigenvectors[c] work1[c]` (Feast.hs:203-208);
  * everything else (ijob 10/11/20/21: factorisations and solves through UMFPACK) is out of scope.

Here the per-column loop is ONE `slc_{d,z}_mulm` over the active range: the column range is chosen by
offsetting the two base pointers, the blocks stay in (pinned) host memory exactly as FEAST owns them,
and the library pipelines upload / product / download by column groups.  `oracle.mulm` restates the
reference loop; tests/test_gpu_boundary.py (test_feast_multiply_work_moving_range) compares every call with it.
"""
import ctypes as C

import numpy as np

from . import _capi
from .matrix import SparseLinearError, _check


class FeastWorkspace:
    """The host-side state `geigSH_` allocates (Feast.hs:149-155): fpm, eigenvectors, work1."""

    def __init__(self, n, m0, dtype=np.complex128, pinned=True):
        self.n, self.m0 = int(n), int(m0)
        self.dtype = np.dtype(dtype)
        self.p = "z" if self.dtype.kind == "c" else "d"
        self.fpm = np.zeros(64, dtype=np.int32)            # feastinit's parameter array (Feast.hs:164)
        self._pinned = []
        self.eigenvectors = self._alloc(pinned)           # MV.replicate (m0 * n) 0
        self.work1 = self._alloc(pinned)
        self.calls = 0

    def _alloc(self, pinned):
        nbytes = max(self.n * self.m0, 1) * self.dtype.itemsize
        if not pinned:
            return np.zeros((self.n, self.m0), dtype=self.dtype, order="F")
        ptr = C.c_void_p()
        _check(_capi.lib().slc_host_alloc(C.byref(ptr), nbytes))
        self._pinned.append(ptr)
        buf = (C.c_char * nbytes).from_address(ptr.value)
        a = np.frombuffer(buf, dtype=self.dtype, count=self.n * self.m0).reshape((self.n, self.m0), order="F")
        a[...] = 0
        return a

    def active_range(self):
        """(ndrop, ntake) as multiplyWork reads them: fpm[23] is 1-based (Feast.hs:204-205)."""
        return int(self.fpm[23]) - 1, int(self.fpm[24])

    def multiplyWork(self, mat):
        """ijob 30 / 40: work1[:, c] <- mat * eigenvectors[:, c] for the active columns, one block product."""
        ndrop, ntake = self.active_range()
        if ndrop < 0 or ntake < 0 or ndrop + ntake > self.m0:
            raise SparseLinearError("multiplyWork: active range [%d, %d) outside 0..%d" % (ndrop, ndrop + ntake, self.m0))
        if mat.ncols != self.n or mat.nrows != self.n:
            raise SparseLinearError("axpy_: column dimension %d does not match operand dimension %d" % (mat.ncols, self.n))
        if mat._p != self.p:
            raise TypeError("multiplyWork: element types differ")
        if ntake == 0:
            return
        el = self.dtype.itemsize
        x = C.c_void_p(self.eigenvectors.ctypes.data + ndrop * self.n * el)   # sliceMat: column c starts at c*n
        y = C.c_void_p(self.work1.ctypes.data + ndrop * self.n * el)
        fn = getattr(_capi.lib(), "slc_%s_mulm" % self.p)
        _check(fn(mat.device().handle, x, max(self.n, 1), y, max(self.n, 1), ntake, None))
        self.calls += 1

    def rci_step(self, ijob, matA, matB):
        """The dispatch of the reference's doM loop (Feast.hs:218-229) for the jobs on this path."""
        if ijob == 30:
            self.multiplyWork(matA)
        elif ijob == 40:
            self.multiplyWork(matB)
        # 10, 11, 20, 21: factorise / solve (UMFPACK) -- not this library's path
        return ijob

    def close(self):
        self.eigenvectors = self.work1 = None
        for p in self._pinned:
            _capi.lib().slc_host_free(p)
        self._pinned = []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def contour_schedule(m0, loops=2):
    """A plausible sequence of (ijob, fpm[23], fpm[24]) as FEAST's RCI issues them for the products:
    per refinement loop the full subspace against A and B (forming Aq, Bq), then, after the reduced
    problem, the products restricted to the columns still inside the search interval (a moving range).
    FEAST decides the real ranges at run time; this only has to exercise every shape of call."""
    out = []
    lo, cnt = 1, m0
    for it in range(loops):
        out.append((30, 1, m0))
        out.append((40, 1, m0))
        lo, cnt = 1 + (it + 1) * max(1, m0 // 8), max(1, m0 - (it + 2) * max(1, m0 // 8))
        if lo + cnt - 1 > m0:
            cnt = m0 - lo + 1
        if cnt > 0:
            out.append((30, lo, cnt))
            out.append((40, lo, cnt))
    return out
