"""BASELINE.json's configs at (or near) full size on the GPU, checked against the oracle and through
size-independent properties (round trips, linearity, structure counts)."""
import numpy as np
import pytest

import oracle
from helpers import assert_same_structure, assert_values_close, to_matrix
from sparse_linear_b200 import generators as g

pytestmark = pytest.mark.gpu


def test_cfg1_laplace2d_1024_spmv(sl):
    gm = g.laplace2d_5pt(1024)
    assert gm.colptr[-1] == 5238784
    m, om = to_matrix(sl, gm), oracle.csc(*gm)
    x = g.uniform_vector(gm.ncols, seed=0)
    got = sl.mulV(m, x)
    np.testing.assert_array_equal(got, oracle.mulv(om, x))         # bit-identical
    assert "sell32" in m.device().plan_info()
    # linearity: A(2x) == 2 A x exactly (scaling by 2 is exact)
    np.testing.assert_array_equal(sl.mulV(m, 2 * x), 2 * got)
    # row sums of the Laplacian: A 1 = boundary indicator >= 0, zero in the interior
    ones = sl.mulV(m, np.ones(gm.ncols))
    assert ones.min() == 0.0 and ones.max() == 2.0


def test_cfg2_complex_spmm_k64(sl):
    gm = g.laplace2d_5pt(1024, np.complex128)
    m, om = to_matrix(sl, gm), oracle.csc(*gm)
    X = g.uniform_block(gm.ncols, 64, seed=0, dtype=np.complex128)
    Y = sl.mulM(m, X)
    # full oracle SpMM on 8 of the 64 columns (bit-identical), SpMV consistency on the rest
    cols = [0, 1, 7, 31, 32, 33, 62, 63]
    want = oracle.mulm(om, X[:, cols])
    np.testing.assert_array_equal(Y[:, cols], want)
    np.testing.assert_array_equal(Y[:, 17], sl.mulV(m, np.ascontiguousarray(X[:, 17])))
    assert_values_close(Y[:, 40], oracle.mulv(om, np.ascontiguousarray(X[:, 40])))


def test_cfg3_spgemm_transpose_assembly_128(sl):
    n = 128
    gm = g.laplace3d_7pt(n)
    assert gm.colptr[-1] == 14581760
    m, om = to_matrix(sl, gm), oracle.csc(*gm)
    c = sl.mm(m, m)
    assert sl.nonZero(c) == 51742208
    want = oracle.mm(om, om)
    assert_same_structure(c, want)
    np.testing.assert_array_equal(c.values, want.values)
    assert int(np.diff(c.pointers).max()) == 25
    t = sl.transpose(m)
    assert t == m                                                      # symmetric: fixed point
    assert sl.hermitian(m)
    rr, cc, vv = g.split_duplicates(gm, seed=0)
    asm = sl.compress(gm.nrows, gm.ncols, rr, cc, vv)
    assert asm == m                                                    # known answer: A bit for bit


def test_cfg4_rmat_scale20_spmv_transpose(sl):
    """cfg 4 at scale 20 (the full scale-24 case runs in bench.py --workload cfg4)."""
    n, r, c, v = g.rmat_triplets(20, edge_factor=16, seed=0)
    m = sl.compress(n, n, r, c, v)
    om = oracle.compress(n, n, r, c, v)
    assert_same_structure(m, om)
    np.testing.assert_array_equal(m.values, om.values)
    x = g.uniform_vector(n, seed=1)
    got = sl.mulV(m, x)
    assert "csr-balanced" in m.device().plan_info()
    assert_values_close(got, oracle.mulv(om, x))
    t = sl.transpose(m)
    ot = oracle.transpose(om)
    assert_same_structure(t, ot)
    np.testing.assert_array_equal(t.values, ot.values)
    assert sl.transpose(t) == m                                        # round trip
    # <A x, y> == <x, A^T y> up to rounding
    y = g.uniform_vector(n, seed=2)
    lhs, rhs = float(np.dot(got, y)), float(np.dot(x, sl.mulV(t, y)))
    assert abs(lhs - rhs) <= 1e-9 * max(1.0, abs(lhs))


def test_cfg5_stencil27_200_spmv(sl):
    gm = g.stencil3d_27pt(200)
    assert gm.colptr[-1] == 213847192
    m, om = to_matrix(sl, gm), oracle.csc(*gm)
    x = g.uniform_vector(gm.ncols, seed=0)
    got = sl.mulV(m, x)
    np.testing.assert_array_equal(got, oracle.mulv(om, x))             # bit-identical at full size
    assert "sell32" in m.device().plan_info()
    # interior rows of this stencil sum to zero: A 1 == 0 there, > 0 on the boundary
    ones = sl.mulV(m, np.ones(gm.ncols)).reshape(200, 200, 200)
    assert np.all(ones[1:-1, 1:-1, 1:-1] == 0.0) and ones.min() >= 0.0 and ones[0, 0, 0] == 19.0
