"""Row-partitioned SpMV kernels on one GPU: all the parts of x live on the same device (a single
process emulates the ranks one after the other), which exercises the partitioned-operand gather."""
import ctypes as C

import numpy as np
import pytest

import oracle
from helpers import random_csc, to_matrix
from sparse_linear_b200 import generators as g
from sparse_linear_b200 import sharded

pytestmark = pytest.mark.gpu


def _dev_array(sl, a):
    lib = sl._capi.lib()
    p = C.c_void_p()
    assert lib.slc_dev_alloc(C.byref(p), max(a.nbytes, 8)) == 0
    assert lib.slc_memcpy_h2d(p, C.c_void_p(a.ctypes.data), a.nbytes) == 0
    return p


@pytest.mark.parametrize("dtype", [np.float64, np.complex128])
@pytest.mark.parametrize("kind", [0, 1, 2, 3])
def test_partitioned_operand_matches_oracle(sl, dtype, kind):
    lib = sl._capi.lib()
    rng = np.random.default_rng(9)
    for gm, nparts in ((g.stencil3d_27pt((8, 6, 7), dtype), 4), (random_csc(rng, 300, 500, 0.05, dtype), 3),
                       (g.laplace2d_5pt(30, dtype), 16)):
        om = oracle.csc(*gm)
        x = rng.uniform(-1, 1, gm.ncols).astype(dtype)
        want = oracle.mulv(om, x)
        xs = sharded.uniform_splits(gm.ncols, nparts).astype(np.int32)
        xparts = [_dev_array(sl, np.ascontiguousarray(x[xs[p]:xs[p + 1]])) for p in range(nparts)]
        table = (C.c_void_p * nparts)(*[p.value for p in xparts])
        rs = sharded.uniform_splits(gm.nrows, nparts)
        fn = getattr(lib, "slc_%s_mulv_dist_dev" % ("z" if np.dtype(dtype).kind == "c" else "d"))
        out = []
        for rank in range(nparts):
            blk = sharded.row_block(gm, rs[rank], rs[rank + 1])
            m = to_matrix(sl, blk)
            m.device().set_spmv_kernel(kind)
            y = np.zeros(blk.nrows, dtype=dtype)
            yd = _dev_array(sl, y)
            rc = fn(m.device().handle, nparts, rank, xs.ctypes.data_as(C.c_void_p), C.cast(table, C.c_void_p), yd, None)
            assert rc == 0, sl._capi.last_error()
            assert lib.slc_memcpy_d2h(C.c_void_p(y.ctypes.data), yd, y.nbytes) == 0
            lib.slc_dev_free(yd)
            out.append(y)
        got = np.concatenate(out)
        if kind == 1:   # row per thread: the reference's summation order
            np.testing.assert_array_equal(got, want)
        else:
            assert np.linalg.norm(got - want) <= 1e-12 * np.linalg.norm(want)
        for p in xparts:
            lib.slc_dev_free(p)


@pytest.mark.parametrize("dtype", [np.float64, np.complex128])
def test_partitioned_block_product_matches_oracle(sl, dtype):
    """slc_*_mulm_dist_dev: X partitioned by row blocks, each part with its own leading dimension;
    every (row, column) is owned by one thread, so the result is the oracle's bit for bit."""
    lib = sl._capi.lib()
    rng = np.random.default_rng(13)
    pfx = "z" if np.dtype(dtype).kind == "c" else "d"
    fn = getattr(lib, "slc_%s_mulm_dist_dev" % pfx)
    for gm, nparts, k in ((g.stencil3d_27pt((8, 6, 7), dtype), 4, 11), (g.laplace2d_5pt(30, dtype), 3, 5),
                          (g.stencil3d_27pt((12, 12, 10), dtype), 2, 64)):
        om = oracle.csc(*gm)
        X = rng.uniform(-1, 1, (gm.ncols, k)).astype(dtype)
        if np.dtype(dtype).kind == "c":
            X = X + 1j * rng.uniform(-1, 1, (gm.ncols, k))
        X = np.asfortranarray(X)
        want = oracle.mulm(om, X)
        xs = sharded.uniform_splits(gm.ncols, nparts).astype(np.int32)
        lds = np.array([int(xs[p + 1] - xs[p]) + 3 * p for p in range(nparts)], dtype=np.int64)   # padded blocks
        parts = []
        for p in range(nparts):
            blk = np.zeros((lds[p], k), dtype=dtype, order="F")
            blk[: xs[p + 1] - xs[p]] = X[xs[p]:xs[p + 1]]
            parts.append(_dev_array(sl, blk))
        table = (C.c_void_p * nparts)(*[p.value for p in parts])
        rs = sharded.uniform_splits(gm.nrows, nparts)
        out = []
        for rank in range(nparts):
            blk = sharded.row_block(gm, rs[rank], rs[rank + 1])
            m = to_matrix(sl, blk)
            Y = np.zeros((blk.nrows, k), dtype=dtype, order="F")
            Yd = _dev_array(sl, Y)
            rc = fn(m.device().handle, nparts, rank, xs.ctypes.data_as(C.c_void_p), C.cast(table, C.c_void_p),
                    lds.ctypes.data_as(C.c_void_p), Yd, blk.nrows, k, None)
            assert rc == 0, sl._capi.last_error()
            assert lib.slc_memcpy_d2h(C.c_void_p(Y.ctypes.data), Yd, Y.nbytes) == 0
            lib.slc_dev_free(Yd)
            out.append(Y)
        np.testing.assert_array_equal(np.concatenate(out, axis=0), want)
        for p in parts:
            lib.slc_dev_free(p)


def test_sharded_wrapper_single_rank(sl):
    gm = g.stencil3d_27pt((10, 9, 8))
    m = to_matrix(sl, gm)
    m.device().set_spmv_kernel(1)    # row per thread: bit-identical to the oracle
    sh = sharded.ShardedSpMV(m, np.array([0, gm.ncols]), 0, 1)
    x = g.uniform_vector(gm.ncols, seed=2)
    sh.set_x(x)
    sh.mulv()
    sl.synchronize()
    np.testing.assert_array_equal(sh.y(), oracle.mulv(oracle.csc(*gm), x))
    sh.close()


def test_bad_partition_is_rejected(sl):
    lib = sl._capi.lib()
    m = to_matrix(sl, g.laplace2d_5pt(4))
    xs = np.array([0, 5, 15], dtype=np.int32)   # does not cover 16 columns
    table = (C.c_void_p * 2)(1, 1)
    rc = lib.slc_d_mulv_dist_dev(m.device().handle, 2, 0, xs.ctypes.data_as(C.c_void_p), C.cast(table, C.c_void_p),
                                 C.c_void_p(1), None)
    assert rc == sl._capi.ERR_DIM and "axpy_: column dimension 16" in sl._capi.last_error()
