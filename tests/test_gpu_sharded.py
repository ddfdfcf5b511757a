"""Row-partitioned SpMV kernels on one GPU: all the parts of x live on the same device (a single
process emulates the ranks one after the other), which exercises the partitioned-operand gather."""
import ctypes as C

import numpy as np
import pytest

import oracle
from helpers import random_csc, rmat_csc, to_matrix
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
                          (g.stencil3d_27pt((12, 12, 10), dtype), 2, 64),
                          (g.stencil3d_27pt((8, 16, 40), dtype), 2, 9)):     # row blocks of 2560 rows: brick plan with a row shift
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


@pytest.mark.parametrize("dtype", [np.float64, np.complex128])
def test_partitioned_operand_compressed_indices(sl, dtype, monkeypatch):
    """The one-launch distributed kernel runs its interior rows on the compressed-index row body
    (one base per (slice, k)); forced on here at test size.  z-slabs of a grid whose lines are whole
    slices, so interior slices compress and boundary slices take the peer path."""
    monkeypatch.setenv("SLC_SELL_KBASE", "2")
    lib = sl._capi.lib()
    rng = np.random.default_rng(10)
    gm = g.stencil3d_27pt((8, 8, 64), dtype)
    nparts, plane = 4, 8 * 64
    om = oracle.csc(*gm)
    x = rng.uniform(-1, 1, gm.ncols).astype(dtype)
    want = oracle.mulv(om, x)
    xs = (sharded.uniform_splits(8, nparts) * plane).astype(np.int32)
    xparts = [_dev_array(sl, np.ascontiguousarray(x[xs[p]:xs[p + 1]])) for p in range(nparts)]
    table = (C.c_void_p * nparts)(*[p.value for p in xparts])
    fn = getattr(lib, "slc_%s_mulv_dist_dev" % ("z" if np.dtype(dtype).kind == "c" else "d"))
    out = []
    for rank in range(nparts):
        blk = sharded.row_block(gm, xs[rank], xs[rank + 1])
        m = to_matrix(sl, blk)
        m.device().set_spmv_kernel(1)
        y = np.zeros(blk.nrows, dtype=dtype)
        yd = _dev_array(sl, y)
        rc = fn(m.device().handle, nparts, rank, xs.ctypes.data_as(C.c_void_p), C.cast(table, C.c_void_p), yd, None)
        assert rc == 0, sl._capi.last_error()
        assert "unit_stride=0.00%" not in m.device().plan_info()
        assert lib.slc_memcpy_d2h(C.c_void_p(y.ctypes.data), yd, y.nbytes) == 0
        lib.slc_dev_free(yd)
        out.append(y)
    np.testing.assert_array_equal(np.concatenate(out), want)
    for p in xparts:
        lib.slc_dev_free(p)


@pytest.mark.parametrize("dtype", [np.float64, np.complex128])
def test_allgather_operand_matches_oracle(sl, dtype):
    """slc_*_mulv_allgather_dev: the slices (odd sizes, so some destinations are not 16-byte aligned)
    are pulled into a full copy of x, then the undistributed kernels run."""
    lib = sl._capi.lib()
    rng = np.random.default_rng(21)
    pfx = "z" if np.dtype(dtype).kind == "c" else "d"
    fn = getattr(lib, "slc_%s_mulv_allgather_dev" % pfx)
    for gm, xs in ((random_csc(rng, 400, 777, 0.03, dtype), [0, 101, 101, 400, 777]),
                   (rmat_csc(10, 8, seed=3, dtype=dtype), None),
                   (g.stencil3d_27pt((7, 6, 5), dtype), [0, 33, 210])):
        om = oracle.csc(*gm)
        x = rng.uniform(-1, 1, gm.ncols).astype(dtype)
        want = oracle.mulv(om, x)
        xs = (sharded.uniform_splits(gm.ncols, 5) if xs is None else np.asarray(xs)).astype(np.int32)
        nparts = len(xs) - 1
        xparts = [_dev_array(sl, np.ascontiguousarray(x[xs[p]:xs[p + 1]])) for p in range(nparts)]
        table = (C.c_void_p * nparts)(*[p.value for p in xparts])
        rownnz = np.bincount(gm.rowidx, minlength=gm.nrows)
        rs = sharded.balanced_splits(rownnz, nparts)
        out = []
        for rank in range(nparts):
            blk = sharded.row_block(gm, rs[rank], rs[rank + 1])
            m = to_matrix(sl, blk)
            m.device().set_spmv_kernel(1)     # row per thread: the reference's summation order
            y = np.zeros(blk.nrows, dtype=dtype)
            yd, xf = _dev_array(sl, y), _dev_array(sl, np.full(gm.ncols, np.nan, dtype=dtype))
            rc = fn(m.device().handle, nparts, rank, xs.ctypes.data_as(C.c_void_p), C.cast(table, C.c_void_p),
                    xf, yd, None)
            assert rc == 0, sl._capi.last_error()
            assert lib.slc_memcpy_d2h(C.c_void_p(y.ctypes.data), yd, y.nbytes) == 0
            lib.slc_dev_free(yd)
            lib.slc_dev_free(xf)
            out.append(y)
        np.testing.assert_array_equal(np.concatenate(out), want)
        for p in xparts:
            lib.slc_dev_free(p)


def test_row_partitioned_spgemm_concatenates_to_the_oracle(sl):
    """C = A B with A cut into row blocks and B replicated: the blocks stacked
    top to bottom are the undistributed product, structure and values bit for bit."""
    rng = np.random.default_rng(5)
    for a, b, world in ((g.laplace3d_7pt(9), g.laplace3d_7pt(9), 4),
                        (random_csc(rng, 120, 90, 0.08), random_csc(rng, 90, 150, 0.06), 3),
                        (rmat_csc(9, 6, seed=1), rmat_csc(9, 6, seed=2), 8)):
        want = oracle.mm(oracle.csc(*a), oracle.csc(*b))
        bm = to_matrix(sl, b)
        blocks = []
        for rank in range(world):
            c, r0, r1 = sharded.mm_row_block(a, bm, rank, world)
            assert c.nrows == r1 - r0 and c.ncols == b.ncols
            blocks.append(g.CSC(c.nrows, c.ncols, c.pointers, c.indices, c.values))
        got = sharded.vstack_blocks(blocks)
        assert got.nrows == a.nrows
        assert np.array_equal(got.colptr, want.colptr) and np.array_equal(got.rowidx, want.rowidx)
        assert np.array_equal(got.values, want.values)


def test_column_partitioned_spgemm_concatenates_to_the_oracle(sl):
    """C = A B with B cut into column blocks and A replicated: the blocks side by side are the
    undistributed product, structure and values bit for bit (uneven splits and an empty block included)."""
    rng = np.random.default_rng(6)
    for a, b, splits in ((g.laplace3d_7pt(9), g.laplace3d_7pt(9), None),
                         (random_csc(rng, 120, 90, 0.08), random_csc(rng, 90, 150, 0.06), [0, 1, 1, 97, 150]),
                         (rmat_csc(9, 6, seed=1), rmat_csc(9, 6, seed=2), None)):
        world = 4 if splits is None else len(splits) - 1
        want = oracle.mm(oracle.csc(*a), oracle.csc(*b))
        am = to_matrix(sl, a)
        blocks = []
        for rank in range(world):
            c, c0, c1 = sharded.mm_col_block(am, b, rank, world, splits)
            assert c.ncols == c1 - c0 and c.nrows == a.nrows
            blocks.append(g.CSC(c.nrows, c.ncols, c.pointers, c.indices, c.values))
        got = sharded.hstack_blocks(blocks)
        assert got.ncols == b.ncols
        assert np.array_equal(got.colptr, want.colptr) and np.array_equal(got.rowidx, want.rowidx)
        assert np.array_equal(got.values, want.values)


@pytest.mark.parametrize("dtype", [np.float64, np.complex128])
def test_sharded_transpose_and_assembly_device_ops(sl, dtype):
    """The ranks of sharded_transpose / sharded_compress run one after the other in this process, the
    local steps on the GPU (slc_*_transpose, slc_*_compress); blocks must equal the oracle's."""
    rng = np.random.default_rng(17)
    m = random_csc(rng, 230, 170, 0.06, dtype)
    world = 3
    cs, rs = sharded.uniform_splits(m.ncols, world), sharded.uniform_splits(m.nrows, world)

    def dev_transpose(blk):
        t = sl.transpose(to_matrix(sl, blk))
        return g.CSC(t.nrows, t.ncols, t.pointers, t.indices, t.values)
    sent = [sharded.transpose_pieces(dev_transpose(sharded.col_block(m, cs[q], cs[q + 1])), rs) for q in range(world)]
    want = oracle.transpose(oracle.csc(*m))
    for p in range(world):
        got = sharded.transpose_combine([sent[q][p] for q in range(world)])
        ref = sharded.col_block(want, rs[p], rs[p + 1])
        assert np.array_equal(got.colptr, ref.colptr) and np.array_equal(got.rowidx, ref.rowidx)
        assert np.array_equal(got.values, ref.values)
    # single rank through the public wrapper (no collective)
    whole = sharded.sharded_transpose(m, np.array([0, m.nrows]))
    assert np.array_equal(whole.colptr, want.colptr) and np.array_equal(whole.rowidx, want.rowidx)

    n = 3000
    r, c = rng.integers(0, 230, n), rng.integers(0, 170, n)
    v = rng.integers(-5, 6, n).astype(dtype)      # integers: sums are exact in any order
    full = oracle.compress(230, 170, r, c, v)
    cuts = [0, 1100, 1900, n]
    sent = [sharded.compress_pieces(r[cuts[q]:cuts[q + 1]], c[cuts[q]:cuts[q + 1]], v[cuts[q]:cuts[q + 1]], cs)
            for q in range(world)]
    for p in range(world):
        got = sharded.compress_combine(230, int(cs[p + 1] - cs[p]), [sent[q][p] for q in range(world)])
        ref = sharded.col_block(full, cs[p], cs[p + 1])
        assert np.array_equal(got.pointers, ref.colptr) and np.array_equal(got.indices, ref.rowidx)
        assert np.array_equal(got.values, ref.values)


def test_sharded_wrapper_allgather_single_rank(sl):
    gm = rmat_csc(9, 8, seed=4)
    m = to_matrix(sl, gm)
    sh = sharded.ShardedSpMV(m, np.array([0, gm.ncols]), 0, 1, mode="allgather")
    x = g.uniform_vector(gm.ncols, seed=2)
    sh.set_x(x)
    sh.mulv()
    sl.synchronize()
    want = oracle.mulv(oracle.csc(*gm), x)
    assert np.linalg.norm(sh.y() - want) <= 1e-12 * np.linalg.norm(want)
    sh.close()


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
