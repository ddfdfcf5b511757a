"""GPU parity: the CUDA path (through the C ABI) against the oracle on the same seeded inputs.

Bar (BASELINE.json north_star): bit-exact sparsity structure for SpGEMM, transpose and assembly;
values within 1e-12 relative norm-wise error; transpose values bit-identical.  Where the kernel
keeps the reference's summation order (row per thread) the test also demands bit-identical values.
"""
import os

import numpy as np
import pytest

import oracle
from helpers import (TOL, arbitrary_matrix, arbitrary_triples, assert_same_structure, assert_values_close,
                     random_csc, rmat_csc, to_matrix)
from sparse_linear_b200 import generators as g

pytestmark = pytest.mark.gpu
DTYPES = [np.float64, np.complex128]
KINDS = {"auto": 0, "sell": 1, "csr-warp": 2, "csr-balanced": 3, "csr-merge": 4}


def vec(rng, n, dtype):
    v = rng.uniform(-1, 1, n)
    if np.dtype(dtype).kind == "c":
        v = v + 1j * rng.uniform(-1, 1, n)
    return v.astype(dtype)


def shapes():
    return [(1, 1, 1.0), (5, 3, 0.5), (33, 65, 0.2), (257, 100, 0.05), (1000, 1000, 0.01), (64, 64, 0.0),
            (300, 7, 0.9), (7, 3000, 0.3)]


# ---- SpMV -----------------------------------------------------------------------------------------

@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("kind", list(KINDS))
def test_mulv_and_axpy_match_oracle(sl, dtype, kind):
    rng = np.random.default_rng(11)
    for nr, nc, dens in shapes():
        om = random_csc(rng, nr, nc, dens, dtype)
        m = to_matrix(sl, om)
        m.device().set_spmv_kernel(KINDS[kind])
        x, y0 = vec(rng, nc, dtype), vec(rng, nr, dtype)
        want = oracle.mulv(om, x)
        got = sl.mulV(m, x)
        assert_values_close(got, want)
        if kind == "sell":   # a thread owns a row and adds in column order: bit-identical
            np.testing.assert_array_equal(got, want)
        want2 = oracle.axpy(om, x, y0)
        got2 = sl.axpy(m, x, y0)
        assert_values_close(got2, want2)
        if kind == "sell":
            np.testing.assert_array_equal(got2, want2)
        y = y0.copy()
        sl.axpy_(m, x, y)
        np.testing.assert_array_equal(y, got2)


@pytest.mark.parametrize("kind", ["csr-warp", "csr-balanced", "csr-merge"])
@pytest.mark.parametrize("dtype", DTYPES)
def test_mulv_irregular_rows(sl, dtype, kind):
    """Power-law-like rows: empty rows, rows spanning several nnz chunks, a dense row."""
    rng = np.random.default_rng(5)
    nr, nc = 3000, 9000
    lens = np.minimum((rng.pareto(1.2, nr) * 3).astype(np.int64), nc)
    lens[7] = nc            # dense row: spans chunks
    lens[8] = 0
    lens[9] = 5000
    lens[2990:] = 0         # trailing empty rows
    rows = np.repeat(np.arange(nr), lens)
    cols = np.concatenate([np.sort(rng.choice(nc, l, replace=False)) for l in lens]) if lens.sum() else []
    vals = vec(rng, len(rows), dtype)
    om = oracle.compress(nr, nc, rows, cols, vals)
    m = to_matrix(sl, om)
    m.device().set_spmv_kernel(KINDS[kind])
    x, y0 = vec(rng, nc, dtype), vec(rng, nr, dtype)
    assert_values_close(sl.mulV(m, x), oracle.mulv(om, x))
    assert_values_close(sl.axpy(m, x, y0), oracle.axpy(om, x, y0))


@pytest.mark.parametrize("kind", ["csr-balanced", "csr-merge"])
@pytest.mark.parametrize("dtype", DTYPES)
def test_mulv_balanced_chunk_edges(sl, dtype, kind):
    """Rows placed against the 2048-nonzero chunk grid of the balanced kernels: a row that is exactly
    one chunk, one that starts on a chunk start and covers three, rows ending exactly on a chunk end
    or on a thread's 8-nonzero range, empty rows on either side of a boundary, a ragged tail."""
    rng = np.random.default_rng(9)
    lens = np.array([2048, 0, 6144, 0, 0, 5, 2043, 8, 8, 16, 1, 7, 0, 2008, 3000, 1096, 0, 4096 + 2048 - 7, 7,
                     1, 1, 1, 0, 13, 4099, 0, 0, 3], dtype=np.int64)
    nr, nc = len(lens) + 2, 7000          # two trailing empty rows
    lens = np.concatenate([lens, [0, 0]])
    rows = np.repeat(np.arange(nr), lens)
    cols = np.concatenate([np.sort(rng.choice(nc, l, replace=False)) for l in lens])
    om = oracle.compress(nr, nc, rows, cols, vec(rng, len(rows), dtype))
    m = to_matrix(sl, om)
    m.device().set_spmv_kernel(KINDS[kind])
    x, y0 = vec(rng, nc, dtype), vec(rng, nr, dtype)
    got = sl.mulV(m, x)
    assert kind in m.device().plan_info()
    assert_values_close(got, oracle.mulv(om, x))
    assert np.all(got[lens == 0] == 0)
    got2 = sl.axpy(m, x, y0)
    assert_values_close(got2, oracle.axpy(om, x, y0))
    np.testing.assert_array_equal(got2[lens == 0], y0[lens == 0])


def test_mulv_summation_order_matches_reference(sl):
    """Same witnesses as tests/test_oracle.py: column order, product and sum rounded separately."""
    m = sl.fromTriples(1, 3, [(0, 0, 1e16), (0, 1, 1.0), (0, 2, -1e16)])
    assert sl.mulV(m, np.ones(3)).tolist() == [0.0]
    a = 1.0 + 2.0 ** -30
    m2 = sl.fromTriples(1, 1, [(0, 0, a)])
    assert sl.axpy(m2, np.array([a]), np.array([-(1.0 + 2.0 ** -29)])).tolist() == [0.0]


@pytest.mark.parametrize("dtype", DTYPES)
def test_mulv_stencils_bit_exact(sl, dtype):
    rng = np.random.default_rng(3)
    for gm in (g.laplace2d_5pt(37, dtype), g.laplace3d_7pt(11, dtype), g.stencil3d_27pt((9, 10, 11), dtype)):
        om = oracle.csc(*gm)
        m = to_matrix(sl, gm)
        x = vec(rng, gm.ncols, dtype)
        got = sl.mulV(m, x)
        assert "sell32" in m.device().plan_info()
        np.testing.assert_array_equal(got, oracle.mulv(om, x))


def test_mulv_identity_and_dimension_errors(sl):
    v = np.arange(17, dtype=np.float64)
    assert np.array_equal(sl.mulV(sl.ident(17), v), v)     # Sparse.hs:41-47
    with pytest.raises(sl.SparseLinearError, match="axpy_: column dimension 17"):
        sl.mulV(sl.ident(17), np.ones(3))
    with pytest.raises(sl.SparseLinearError, match="axpy_: row dimension 17"):
        sl.axpy(sl.ident(17), v, np.ones(3))


# ---- SpMM -----------------------------------------------------------------------------------------

@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("kind", ["auto", "csr-balanced"])
def test_mulm_matches_oracle(sl, dtype, kind):
    rng = np.random.default_rng(21)
    for (nr, nc, dens), k in zip(shapes(), [1, 3, 4, 9, 64, 5, 8, 2]):
        om = random_csc(rng, nr, nc, dens, dtype)
        m = to_matrix(sl, om)
        m.device().set_spmv_kernel(KINDS[kind])
        X = np.asfortranarray(np.stack([vec(rng, nc, dtype) for _ in range(k)], axis=1))
        want = oracle.mulm(om, X)
        got = sl.mulM(m, X)
        assert got.shape == (nr, k)
        np.testing.assert_array_equal(got, want)     # a thread owns (row, column): reference order


@pytest.mark.parametrize("dtype", DTYPES)
def test_mulv_sell_compressed_indices(sl, dtype):
    """Sliced-ELL plans replace the 32 indices of a (slice, k) by one base when they are base + lane
    (stencils away from grid boundaries).  Same rows, same order: bit-identical to the oracle, with
    and without the compression, for grids whose lines do and do not align with slices, a band
    matrix whose every slice qualifies, and axpy."""
    rng = np.random.default_rng(55)
    for gm in (g.stencil3d_27pt((9, 10, 11), dtype), g.stencil3d_27pt((4, 32, 64), dtype), g.laplace2d_5pt(37, dtype),
               g.tridiag(1000, dtype=dtype), g.laplace3d_7pt(12, dtype)):
        om = oracle.csc(*gm)
        x, y0 = vec(rng, gm.ncols, dtype), vec(rng, gm.nrows, dtype)
        want, want_acc = oracle.mulv(om, x), oracle.axpy(om, x, y0)
        infos = []
        for flag in ("2", "0"):      # 2 = compress whatever the size
            os.environ["SLC_SELL_KBASE"] = flag
            try:
                m = to_matrix(sl, gm)
                m.device().set_spmv_kernel(KINDS["sell"])
                np.testing.assert_array_equal(sl.mulV(m, x), want)
                np.testing.assert_array_equal(sl.axpy(m, x, y0), want_acc)
                infos.append(m.device().plan_info())
            finally:
                del os.environ["SLC_SELL_KBASE"]
        assert "unit_stride=0.00%" in infos[1]
        if gm.nrows in (1000, 4 * 32 * 64):   # the band matrix and the grid whose lines are whole slices compress
            assert "unit_stride=0.00%" not in infos[0], infos[0]


@pytest.mark.parametrize("dtype", DTYPES)
def test_mulv_balanced_hot_columns(sl, dtype):
    """Power-law matrices on the balanced plan keep x at their densest columns in shared memory
    (k_spmv_bal_hot, resident blocks walking the chunks).  Small tables force hot and cold gathers in
    the same warp, a table larger than the chunk count exercises idle groups, axpy exercises ACC."""
    rng = np.random.default_rng(41)
    for scale, ef, cap in ((12, 8, "64"), (14, 6, "1000"), (10, 40, "7"), (15, 3, "4096")):
        om = rmat_csc(scale, ef, seed=scale, dtype=dtype)
        x = vec(rng, om.ncols, dtype)
        y0 = vec(rng, om.nrows, dtype)
        want, want_acc = oracle.mulv(om, x), oracle.axpy(om, x, y0)
        os.environ["SLC_BAL_HOT_K"] = cap
        try:
            m = to_matrix(sl, om)
            m.device().set_spmv_kernel(KINDS["csr-balanced"])
            got, got_acc = sl.mulV(m, x), sl.axpy(m, x, y0)
        finally:
            del os.environ["SLC_BAL_HOT_K"]
        assert_values_close(got, want)
        assert_values_close(got_acc, want_acc)
        # the table changes where x is read from, not the arithmetic: same bits as the plain kernel
        os.environ["SLC_BAL_HOT"] = "0"
        try:
            m2 = to_matrix(sl, om)
            m2.device().set_spmv_kernel(KINDS["csr-balanced"])
            np.testing.assert_array_equal(sl.mulV(m2, x), got)
        finally:
            del os.environ["SLC_BAL_HOT"]


@pytest.mark.parametrize("dtype", DTYPES)
def test_mulm_sell_plan_all_row_lengths(sl, dtype):
    """The SELL-plan SpMM kernels (rows in registers up to 8 entries, rows staged in shared memory
    above that, entries past the shared-memory capacity re-read from HBM): ragged rows, empty rows,
    column counts that are not a multiple of the column tile."""
    rng = np.random.default_rng(77)
    for nr, nc, maxlen, k in [(70, 90, 3, 5), (300, 300, 8, 7), (500, 400, 13, 9), (257, 1000, 27, 17),
                              (129, 3000, 150, 6), (64, 64, 40, 64)]:
        lens = rng.integers(0, maxlen + 1, nr)
        lens[rng.integers(0, nr)] = maxlen
        lens = np.minimum(lens, nc)
        rows = np.repeat(np.arange(nr), lens)
        cols = np.concatenate([np.sort(rng.choice(nc, l, replace=False)) for l in lens])
        om = oracle.compress(nr, nc, rows, cols, vec(rng, len(rows), dtype))
        m = to_matrix(sl, om)
        m.device().set_spmv_kernel(KINDS["sell"])
        X = np.asfortranarray(np.stack([vec(rng, nc, dtype) for _ in range(k)], axis=1))
        got = sl.mulM(m, X)
        assert "sell32" in m.device().plan_info()
        np.testing.assert_array_equal(got, oracle.mulm(om, X))


@pytest.mark.parametrize("dtype", DTYPES)
def test_mulm_stencils_bit_exact(sl, dtype):
    rng = np.random.default_rng(4)
    for gm, k in ((g.laplace3d_7pt(9, dtype), 10), (g.stencil3d_27pt((9, 10, 11), dtype), 19),
                  (g.stencil3d_27pt((12, 7, 8), dtype), 64)):
        om, m = oracle.csc(*gm), to_matrix(sl, gm)
        X = np.asfortranarray(np.stack([vec(rng, gm.ncols, dtype) for _ in range(k)], axis=1))
        np.testing.assert_array_equal(sl.mulM(m, X), oracle.mulm(om, X))


@pytest.mark.parametrize("cluster", ["1", "0"])
@pytest.mark.parametrize("dtype", DTYPES)
def test_mulm_banded_window_kernel(sl, dtype, cluster, monkeypatch):
    """Banded matrices with short rows take the TMA-staged kernels: k_spmm_window (one CTA per 1024-row
    block) and k_spmm_window2 (two CTAs per cluster sharing the union of their windows through TMA
    multicast).  Sizes around one and several blocks, an odd number of blocks (the last cluster has a CTA
    with no rows), a ragged last block, fewer columns than pipeline stages, many more columns than stages,
    and a band wide enough that the stage count drops to the minimum."""
    monkeypatch.setenv("SLC_SPMM_CLUSTER", cluster)
    rng = np.random.default_rng(31)
    cases = [(g.laplace2d_5pt(32, dtype), 3), (g.laplace2d_5pt(46, dtype), 9), (g.laplace2d_5pt(90, dtype), 20),
             (g.tridiag(5000, dtype=dtype), 1), (g.tridiag(5000, dtype=dtype), 33), (g.laplace2d_5pt(64, dtype), 17),
             (g.laplace2d_5pt(300, dtype), 2)]
    for gm, k in cases:
        om, m = oracle.csc(*gm), to_matrix(sl, gm)
        X = np.asfortranarray(np.stack([vec(rng, gm.ncols, dtype) for _ in range(k)], axis=1))
        np.testing.assert_array_equal(sl.mulM(m, X), oracle.mulm(om, X))
    assert sl.spmm_window_error() == 0


def test_mulm_reports_a_staged_kernel_that_gave_up(sl, monkeypatch):
    """A TMA-staged block product whose bounded barrier wait gives up leaves a stale block; the host-operand
    product must say so (SLC_ERR_CUDA) instead of returning SLC_OK, and the next product must start clean.
    A real give-up needs a protocol bug, so the word is raised from the host (SLC_DEBUG_WINDOW_FAULT)."""
    gm = g.laplace2d_5pt(46)
    om, m = oracle.csc(*gm), to_matrix(sl, gm)
    X = g.uniform_block(gm.ncols, 5, seed=3)
    np.testing.assert_array_equal(sl.mulM(m, X), oracle.mulm(om, X))
    monkeypatch.setenv("SLC_DEBUG_WINDOW_FAULT", "1")
    with pytest.raises(sl.SparseLinearError, match="gave up waiting on a shared-memory barrier"):
        sl.mulM(m, X)
    monkeypatch.delenv("SLC_DEBUG_WINDOW_FAULT")
    assert sl.spmm_window_error() == 0                     # reported once, then cleared
    np.testing.assert_array_equal(sl.mulM(m, X), oracle.mulm(om, X))
    from sparse_linear_b200 import sharded
    sm = sharded.ShardedMatrix(to_matrix(sl, g.stencil3d_27pt((6, 8, 64))), 1)
    Xs = g.uniform_block(6 * 8 * 64, 3, seed=4)
    monkeypatch.setenv("SLC_DEBUG_WINDOW_FAULT", "2")
    with pytest.raises(sl.SparseLinearError, match="wider than its plan"):
        sm.mulM(Xs)
    monkeypatch.delenv("SLC_DEBUG_WINDOW_FAULT")
    np.testing.assert_array_equal(sm.mulM(Xs), oracle.mulm(oracle.csc(*g.stencil3d_27pt((6, 8, 64))), Xs))
    sm.close()


@pytest.mark.parametrize("dtype", DTYPES)
def test_mulm_panel_line_tiling(sl, dtype):
    """Stencils large enough that k_spmm_panel tiles its blocks over grid lines (the slice-order
    table): line strides that are and are not multiples of 32, a ragged last slice, a 2-D 9-point
    grid (one level of lines only).  Any slice order must give the reference's rows bit for bit."""
    rng = np.random.default_rng(8)
    nine = g._stencil((70, 90), [(dy, dx) for dy in (-1, 0, 1) for dx in (-1, 0, 1)],
                      [8.0 if (dy, dx) == (0, 0) else -1.0 for dy in (-1, 0, 1) for dx in (-1, 0, 1)], dtype)
    for gm, k in ((g.stencil3d_27pt((12, 40, 48), dtype), 9), (g.stencil3d_27pt((10, 30, 50), dtype), 16),
                  (g.stencil3d_27pt((7, 19, 33), dtype), 3), (nine, 8)):
        om = oracle.csc(*gm)
        X = np.asfortranarray(np.stack([vec(rng, gm.ncols, dtype) for _ in range(k)], axis=1))
        want = oracle.mulm(om, X)
        for stride, width in (("-1", "1"), ("-1", "2"), ("64", "1"), ("0", "1")):
            os.environ["SLC_PANEL_STRIDE"], os.environ["SLC_PANEL_WIDTH"] = stride, width
            try:
                np.testing.assert_array_equal(sl.mulM(to_matrix(sl, gm), X), want)
            finally:
                del os.environ["SLC_PANEL_STRIDE"], os.environ["SLC_PANEL_WIDTH"]


def test_mulm_stencil_block_and_column_range(sl):
    """FEAST's multiplyWork: only an active column range is multiplied (Feast.hs:203-208)."""
    rng = np.random.default_rng(2)
    gm = g.laplace2d_5pt(20, np.complex128)
    om, m = oracle.csc(*gm), to_matrix(sl, gm)
    X = np.asfortranarray(rng.uniform(-1, 1, (400, 12)) + 1j * rng.uniform(-1, 1, (400, 12)))
    full = sl.mulM(m, X)
    np.testing.assert_array_equal(full, oracle.mulm(om, X))
    part = sl.mulM(m, X[:, 3:8])
    np.testing.assert_array_equal(part, full[:, 3:8])
    with pytest.raises(sl.SparseLinearError, match="mulM: inner dimension mismatch"):
        sl.mulM(m, X[:10])


# ---- transpose ---------------------------------------------------------------------------------------

@pytest.mark.parametrize("bucket", ["1", "0"])
@pytest.mark.parametrize("dtype", DTYPES)
def test_transpose_and_ctrans(sl, dtype, bucket, monkeypatch):
    """Both transpose paths: count / place / sort-per-row (rows of at most 512 entries; (40, 2000) makes a
    block take several passes over its rows, (5000, 300) spreads over many blocks, (7, 3000) has rows too
    long for it and falls through) and the radix sort."""
    monkeypatch.setenv("SLC_TR_BUCKET", bucket)
    rng = np.random.default_rng(31)
    for nr, nc, dens in shapes() + [(40, 2000, 0.2), (5000, 300, 0.03)]:
        om = random_csc(rng, nr, nc, dens, dtype)
        m = to_matrix(sl, om)
        for fn, ofn in ((sl.transpose, oracle.transpose), (sl.ctrans, oracle.ctrans)):
            got, want = fn(m), ofn(om)
            assert_same_structure(got, want)
            np.testing.assert_array_equal(got.values, want.values)   # moved bit for bit
        tt = sl.transpose(sl.transpose(m))
        assert tt == m


@pytest.mark.parametrize("dtype", DTYPES)
def test_transpose_long_runs_of_empty_rows(sl, dtype):
    """Row pointers over long runs of empty rows (a column block of a banded matrix: entries near the top and the
    bottom only, nothing in between, nothing after the last used row) are filled by k_ptr_fill_gaps."""
    rng = np.random.default_rng(77)
    nr, nc = 9000, 400
    rows = np.concatenate([rng.integers(0, 20, 600), rng.integers(4000, 4003, 50), rng.integers(8100, 8130, 600)])
    cols = rng.integers(0, nc, len(rows))
    om = oracle.compress(nr, nc, rows, cols, vec(rng, len(rows), dtype))
    m = to_matrix(sl, om)
    got, want = sl.transpose(m), oracle.transpose(om)
    assert_same_structure(got, want)
    np.testing.assert_array_equal(got.values, want.values)
    assert sl.transpose(got) == m
    x = vec(rng, nc, dtype)
    np.testing.assert_array_equal(sl.mulV(m, x), oracle.mulv(om, x))     # the row-major mirror takes the same path


def test_ctrans_known_answers_and_hermitian(sl):
    """sparse-linear/tests/Sparse.hs:61-73"""
    import json, os
    cases = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "reference_known_answers.json")))
    for c in cases["ctrans_fixed_points"]:
        m = sl.fromTriples(c["nrows"], c["ncols"], [(r, col, complex(re, im)) for r, col, re, im in c["triples"]],
                           dtype=np.complex128)
        assert sl.ctrans(m) == m, c["name"]
        assert sl.hermitian(m)
    rng = np.random.default_rng(1)
    a = to_matrix(sl, random_csc(rng, 40, 40, 0.2, np.complex128))
    assert not sl.hermitian(a)
    assert sl.hermitian(a + sl.ctrans(a))
    assert not sl.hermitian(to_matrix(sl, random_csc(rng, 4, 5, 0.5, np.complex128)))
    d = sl.diag(np.arange(9, dtype=np.float64))
    assert sl.transpose(d) == d                                       # Sparse.hs:56-59


# ---- triplet assembly ----------------------------------------------------------------------------------

@pytest.mark.parametrize("bucket", ["1", "coarse", "0"])
@pytest.mark.parametrize("dtype", DTYPES)
def test_compress_matches_oracle(sl, dtype, bucket, monkeypatch):
    """The assembly paths: count / place / sort-per-column (short columns; (513, 129) makes a block
    take several passes over its span, (64, 3000) spreads over many blocks, (2000, 3) has columns
    longer than the bucket limit and falls through to the sort), the same behind the coarse column
    partition (forced on: by default only inputs that outgrow L2 take it), and the full radix sort."""
    monkeypatch.setenv("SLC_ASM_BUCKET", "0" if bucket == "0" else "1")
    monkeypatch.setenv("SLC_ASM_COARSE", "1" if bucket == "coarse" else "0")
    rng = np.random.default_rng(41)
    for nr, nc in [(1, 1), (3, 5), (40, 40), (200, 10), (10, 2000), (513, 129), (64, 3000), (2000, 3)]:
        for integer in (True, False):
            r, c, v = arbitrary_triples(rng, nr, nc, dtype, integer)
            want = oracle.compress(nr, nc, r, c, v)
            got = sl.compress(nr, nc, r, c, v)
            assert_same_structure(got, want)
            np.testing.assert_array_equal(got.values, want.values)    # duplicates summed in input order
            assert oracle.check_format(oracle.csc(got.nrows, got.ncols, got.pointers, got.indices, got.values)) == 0


def test_compress_many_duplicates_zeros_and_empty(sl):
    rng = np.random.default_rng(4)
    r = rng.integers(0, 3, 5000)
    c = rng.integers(0, 3, 5000)
    v = rng.uniform(-1, 1, 5000)
    v[::7] = 0.0
    want, got = oracle.compress(3, 3, r, c, v), sl.compress(3, 3, r, c, v)
    assert_same_structure(got, want)
    np.testing.assert_array_equal(got.values, want.values)
    e = sl.compress(4, 6, [], [], np.zeros(0))
    assert e.pointers.tolist() == [0] * 7 and sl.nonZero(e) == 0
    z = sl.fromTriples(2, 2, [(0, 0, 0.0), (1, 1, 1.0), (1, 1, -1.0)])
    assert z.pointers.tolist() == [0, 1, 2] and z.values.tolist() == [0.0, 0.0]   # zeros are kept


@pytest.mark.parametrize("coarse", ["0", "1"])
@pytest.mark.parametrize("dtype", DTYPES)
def test_compress_any_duplicate_order_stays_inside_the_tolerance(sl, dtype, coarse, monkeypatch):
    """The reference sorts the triplets with an unstable introsort before it sums duplicates
    (Sparse.hs:219,257-280), so the order in which the terms of one entry are added is unspecified; the
    oracle and the device both add them in input order.  Feeding the same triplets in several random
    orders covers whatever order the reference might have used: the structure must not change at all, and
    the values must stay within north_star's 1e-12 norm-wise bound of the oracle on the original order."""
    monkeypatch.setenv("SLC_ASM_COARSE", coarse)
    rng = np.random.default_rng(17)
    nr, nc, n = 300, 200, 120_000                      # two duplicates per entry on average, up to ~10
    r = rng.integers(0, nr, n)
    c = rng.integers(0, nc, n)
    v = rng.uniform(-1, 1, n) * 10.0 ** rng.integers(-6, 7, n)     # mixed magnitudes: order matters in the last bits
    v = v.astype(dtype)
    if np.dtype(dtype).kind == "c":
        v = v + 1j * rng.uniform(-1, 1, n) * 10.0 ** rng.integers(-6, 7, n)
    want = oracle.compress(nr, nc, r, c, v)
    differing = 0
    for trial in range(6):
        perm = rng.permutation(n)
        got = sl.compress(nr, nc, r[perm], c[perm], v[perm])
        assert_same_structure(got, want)
        assert_values_close(got.values, want.values)
        # and the device agrees bit for bit with the oracle fed the same order
        np.testing.assert_array_equal(got.values, oracle.compress(nr, nc, r[perm], c[perm], v[perm]).values)
        differing += int(not np.array_equal(got.values, want.values))
    assert differing > 0          # the orders did change some last bits: the test is not vacuous


@pytest.mark.parametrize("coarse", ["0", "1"])
def test_compress_bounds_errors(sl, coarse, monkeypatch):
    """Sparse.hs:196-212: rows are checked before columns; the first offending position is reported."""
    monkeypatch.setenv("SLC_ASM_COARSE", coarse)
    with pytest.raises(sl.SparseLinearError, match=r"compress: row index out of bounds \(0,3\) at 1") as e:
        sl.compress(3, 3, [0, 3, 5], [0, 9, 0], [1.0, 1.0, 1.0])
    assert e.value.pos == 1 and e.value.code == sl._capi.ERR_ROW_BOUNDS
    with pytest.raises(sl.SparseLinearError, match=r"compress: column index out of bounds \(0,3\) at 1") as e:
        sl.compress(3, 3, [0, 1, 2], [0, 9, -1], [1.0, 1.0, 1.0])
    assert e.value.pos == 1 and e.value.code == sl._capi.ERR_COL_BOUNDS
    n = 100000
    r = np.zeros(n, dtype=np.int64)
    r[[77777, 99999]] = -1
    with pytest.raises(sl.SparseLinearError) as e:
        sl.compress(5, 5, r, np.zeros(n, dtype=np.int64), np.ones(n))
    assert e.value.pos == 77777


# ---- SpGEMM ---------------------------------------------------------------------------------------------

@pytest.mark.parametrize("merge", ["1", "0"])     # thread-per-column merge classes on / hash tables only
@pytest.mark.parametrize("dtype", DTYPES)
def test_mm_matches_oracle(sl, dtype, merge, monkeypatch):
    monkeypatch.setenv("SLC_MM_MERGE", merge)
    rng = np.random.default_rng(51)
    cases = [(1, 1, 1, 1.0, 1.0), (6, 4, 5, 0.5, 0.5), (50, 60, 70, 0.1, 0.1), (300, 300, 300, 0.02, 0.02),
             (500, 400, 300, 0.05, 0.005), (2000, 64, 900, 0.03, 0.08),   # <= 4 and <= 8 entries per column of B
             (40, 40, 40, 0.0, 0.3), (64, 512, 64, 0.5, 0.5),      # class 1/2 columns
             (3000, 200, 50, 0.6, 0.9)]                            # > 2048 products: dense fallback
    for nr, inner, nc, da, db in cases:
        oa, ob = random_csc(rng, nr, inner, da, dtype), random_csc(rng, inner, nc, db, dtype)
        want = oracle.mm(oa, ob)
        got = sl.mm(to_matrix(sl, oa), to_matrix(sl, ob))
        assert_same_structure(got, want)
        np.testing.assert_array_equal(got.values, want.values)    # terms added in increasing inner index


def test_mm_structural_zeros_and_errors(sl):
    a = sl.fromTriples(1, 2, [(0, 0, 1.0), (0, 1, 1.0)])
    b = sl.fromTriples(2, 1, [(0, 0, 1.0), (1, 0, -1.0)])
    c = a * b
    assert c.pointers.tolist() == [0, 1] and c.indices.tolist() == [0] and c.values.tolist() == [0.0]
    z = sl.fromTriples(2, 2, [(0, 0, 0.0), (1, 1, 5.0)])
    assert (z * z).pointers.tolist() == [0, 1, 2]
    # the accumulator starts from +0 (ScatterGather reset), so a lone -0 term comes out as +0
    n = sl.fromTriples(1, 1, [(0, 0, -1.0)]) * sl.fromTriples(1, 1, [(0, 0, 0.0)])
    assert n.values.tolist() == [0.0] and not np.signbit(n.values[0])
    with pytest.raises(sl.SparseLinearError, match="mm: inner dimension mismatch"):
        sl.mm(sl.ident(3), sl.ident(4))


@pytest.mark.parametrize("merge", ["1", "0"])
def test_mm_stencil_squared(sl, merge, monkeypatch):
    monkeypatch.setenv("SLC_MM_MERGE", merge)
    gm = g.laplace3d_7pt(12)
    om, m = oracle.csc(*gm), to_matrix(sl, gm)
    want, got = oracle.mm(om, om), sl.mm(m, m)
    assert_same_structure(got, want)
    np.testing.assert_array_equal(got.values, want.values)
    assert int(np.diff(got.pointers).max()) == 25


# ---- lin / kronecker --------------------------------------------------------------------------------------

@pytest.mark.parametrize("dtype", DTYPES)
def test_lin_and_kronecker_match_oracle(sl, dtype):
    rng = np.random.default_rng(61)
    for nr, nc in [(1, 1), (9, 4), (80, 90), (400, 3)]:
        oa, ob = arbitrary_matrix(rng, nr, nc, dtype, False), arbitrary_matrix(rng, nr, nc, dtype, False)
        al, be = (0.5 - 2j, 3.0 + 1j) if np.dtype(dtype).kind == "c" else (0.5, -3.0)
        want = oracle.lin(al, oa, be, ob)
        got = sl.lin(al, to_matrix(sl, oa), be, to_matrix(sl, ob))
        assert_same_structure(got, want)
        np.testing.assert_array_equal(got.values, want.values)
    for (ar, ac), (br, bc) in [((1, 1), (1, 1)), ((3, 4), (5, 2)), ((12, 9), (7, 11))]:
        oa, ob = arbitrary_matrix(rng, ar, ac, dtype, False), arbitrary_matrix(rng, br, bc, dtype, False)
        want = oracle.kronecker(oa, ob)
        got = sl.kronecker(to_matrix(sl, oa), to_matrix(sl, ob))
        assert_same_structure(got, want)
        np.testing.assert_array_equal(got.values, want.values)


def test_configs_built_on_device_equal_generators(sl):
    """cfg 1 and cfg 3 as the configs describe them: kronecker sums, built by the device kernels."""
    n = 9
    T, I = to_matrix(sl, g.tridiag(n)), sl.ident(n)
    k = sl.kronecker
    assert (k(I, T) + k(T, I)) == to_matrix(sl, g.laplace2d_5pt(n))
    assert (k(k(I, I), T) + k(k(I, T), I) + k(k(T, I), I)) == to_matrix(sl, g.laplace3d_7pt(n))


# ---- marshalling round trip -----------------------------------------------------------------------------------

def test_upload_download_round_trip_and_validation(sl):
    """fromForeign . withConstMatrix == id (Sparse.hs:138-145)"""
    rng = np.random.default_rng(71)
    for dtype in DTYPES:
        om = random_csc(rng, 50, 60, 0.1, dtype)
        m = to_matrix(sl, om)
        back = sl.Matrix._from_devmatrix(m.device())    # downloads what was uploaded
        assert back == m
    bad = sl.Matrix(2, 3, [0, 2, 3], [2, 1, 0], np.ones(3))      # rows not increasing in column 0
    with pytest.raises(sl.SparseLinearError, match="upload"):
        bad.device()
    bad2 = sl.Matrix(2, 3, [0, 1, 2], [0, 7], np.ones(2))        # row out of range
    with pytest.raises(sl.SparseLinearError, match="upload"):
        bad2.device()


# ---- the register-blocked brick SpMM for 27-offset matrices (csrc/slc_spmm_brick.cu) ---------------------

def _band27(n, L, P, rng, keep=1.0):
    """Every row r holds the 27 offsets a P + b L + c that stay inside [0, n): unlike a grid stencil, rows at
    'line ends' keep their wrapped neighbours, so the kernel's linear addressing (not a grid assumption) is
    what makes it right.  keep < 1 drops random entries (rows with holes: the masked path)."""
    offs = np.array([a * P + b * L + c for a in (-1, 0, 1) for b in (-1, 0, 1) for c in (-1, 0, 1)])
    cols = np.arange(n)[:, None] + offs[None, :]           # row r, slot s -> column
    ok = (cols >= 0) & (cols < n)
    if keep < 1.0:
        ok &= rng.random(ok.shape) < keep
        ok[n // 2] = (cols[n // 2] >= 0) & (cols[n // 2] < n)      # the row the detector reads stays complete
    r, s = np.nonzero(ok)
    return oracle.compress(n, n, r, cols[r, s], rng.uniform(-1, 1, len(r)))


@pytest.mark.parametrize("jt,rows", [("1", "512"), ("2", "512"), ("4", "512"), ("1", "256"), ("2", "256")])
def test_spmm_brick_matches_oracle(sl, monkeypatch, jt, rows):
    """The TMA-staged kernel for 27-offset matrices (k_spmm_planes): bit-exact against the oracle for every
    block shape, with odd and even strides, wrapped neighbours, holes (masked path), column counts that leave
    a partial tile, and the accumulate form."""
    monkeypatch.setenv("SLC_PLANES_JT", jt)
    monkeypatch.setenv("SLC_PLANES_ROWS", rows)
    rng = np.random.default_rng(41)
    # (matrix, must the staged kernel serve it).  Grids are big enough for a sliced plan (padding <= 12.5 %).
    cases = [(g.stencil3d_27pt((14, 30, 67)), True),        # odd line length: window starts that need rounding
             (g.stencil3d_27pt((16, 32, 64)), True),        # even strides
             (_band27(6000, 50, 400, rng), True),           # wrapped neighbours present
             (_band27(7000, 40, 400, rng, keep=0.97), True),   # a few holes, still a sliced plan: masked path
             (g.stencil3d_27pt((9, 10, 37)), True),         # small grid (5 % padding: still a sliced plan), blocks mostly boundary
             (_band27(5000, 24, 240, rng, keep=0.8), False)]   # many holes: the same
    for gm, staged in cases:
        m, om = to_matrix(sl, gm), oracle.csc(*gm)
        for k in (1, 4, 7, 64):
            X = g.uniform_block(gm.ncols, k, seed=k)
            got = sl.mulM(m, X)
            plan = m.device().plan_info()
            assert ("brick=1" in plan and plan.startswith("sell32")) == staged, plan
            np.testing.assert_array_equal(got, oracle.mulm(om, X))
        Y0 = g.uniform_block(gm.nrows, 5, seed=3)
        X = g.uniform_block(gm.ncols, 5, seed=4)
        want = np.stack([oracle.axpy(om, np.ascontiguousarray(X[:, j]), np.ascontiguousarray(Y0[:, j])) for j in range(5)], axis=1)
        np.testing.assert_array_equal(sl.mulM(m, X, Y0), want)
    assert sl.spmm_window_error() == 0     # no bounded barrier wait of the staged kernels gave up


def test_spmm_brick_switch_gives_the_same_bits(sl, monkeypatch):
    """SLC_SPMM_BRICK=0 (the panel kernel) and the staged kernel agree bit for bit."""
    gm = g.stencil3d_27pt((12, 20, 50))
    m = to_matrix(sl, gm)
    X = g.uniform_block(gm.ncols, 9, seed=2)
    a = sl.mulM(m, X)
    assert "brick=1" in m.device().plan_info()
    monkeypatch.setenv("SLC_SPMM_BRICK", "0")
    b = sl.mulM(m, X)
    np.testing.assert_array_equal(a, b)


def test_spmm_brick_vetoed_by_an_off_pattern_entry(sl):
    """One entry outside the 27 offsets (far away, in a row the detector does not sample): the plan builder
    must notice and the general kernel must serve the product."""
    gm = g.stencil3d_27pt((8, 12, 40))
    rows = np.repeat(np.arange(gm.ncols), np.diff(gm.colptr))      # symmetric pattern: (col, row) listing is fine
    r = np.concatenate([gm.rowidx, [7]])
    c = np.concatenate([rows, [3000]])
    v = np.concatenate([gm.values, [0.5]])
    bad = oracle.compress(gm.nrows, gm.ncols, r, c, v)
    m, om = to_matrix(sl, bad), oracle.csc(*bad)
    X = g.uniform_block(bad.ncols, 6, seed=1)
    np.testing.assert_array_equal(sl.mulM(m, X), oracle.mulm(om, X))
    assert "brick=0" in m.device().plan_info()
