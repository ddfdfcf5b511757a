"""The config builders give exactly the matrices the configs name (kronecker sums), checked with
the oracle's kronecker / lin (Data/Matrix/Sparse.hs:597-638, 401-431).  CPU only."""
import numpy as np

import oracle
from sparse_linear_b200 import generators as g


def as_orc(m):
    return oracle.csc(m.nrows, m.ncols, m.colptr, m.rowidx, m.values)


def same(a, b):
    return (a.nrows == b.nrows and a.ncols == b.ncols and np.array_equal(a.colptr, b.colptr)
            and np.array_equal(a.rowidx, b.rowidx) and np.array_equal(a.values, b.values))


def test_laplace2d_is_kronecker_sum():
    n = 7
    T, I = as_orc(g.tridiag(n)), oracle.ident(n)
    want = oracle.add(oracle.kronecker(I, T), oracle.kronecker(T, I))
    assert same(as_orc(g.laplace2d_5pt(n)), want)


def test_laplace3d_is_kronecker_sum():
    n = 5
    T, I = as_orc(g.tridiag(n)), oracle.ident(n)
    k = oracle.kronecker
    want = oracle.add(oracle.add(k(k(I, I), T), k(k(I, T), I)), k(k(T, I), I))
    assert same(as_orc(g.laplace3d_7pt(n)), want)


def test_stencil27_pattern_is_triple_kronecker():
    n = 4
    P = as_orc(g.tridiag(n, 1.0, 1.0, 1.0))
    pat = oracle.kronecker(oracle.kronecker(P, P), P)
    m = as_orc(g.stencil3d_27pt(n))
    assert np.array_equal(m.colptr, pat.colptr) and np.array_equal(m.rowidx, pat.rowidx)
    cols = oracle.decompress(m.colptr)
    assert np.all(m.values[m.rowidx == cols] == 26.0) and np.all(m.values[m.rowidx != cols] == -1.0)
    assert oracle.hermitian(m)


def test_closed_form_counts():
    assert g.laplace2d_5pt(64).colptr[-1] == 5 * 64 * 64 - 4 * 64
    assert g.laplace3d_7pt(16).colptr[-1] == 7 * 16 ** 3 - 6 * 16 ** 2
    assert g.stencil3d_27pt(12).colptr[-1] == (3 * 12 - 2) ** 3
    assert g.stencil3d_27pt((3, 4, 5)).colptr[-1] == (3 * 3 - 2) * (3 * 4 - 2) * (3 * 5 - 2)


def test_rmat_and_split_duplicates():
    n, r, c, v = g.rmat_triplets(8, edge_factor=4, seed=3)
    assert n == 256 and len(r) == 1024 and r.min() >= 0 and r.max() < n and c.max() < n
    n2, r2, c2, v2 = g.rmat_triplets(8, edge_factor=4, seed=3)
    assert np.array_equal(r, r2) and np.array_equal(v, v2)
    m = g.laplace3d_7pt(4)
    rr, cc, vv = g.split_duplicates(m, seed=1)
    back = oracle.compress(m.nrows, m.ncols, rr, cc, vv)
    assert same(back, as_orc(m))   # 0.25 v + 0.75 v == v exactly, in either order


def test_algorithmic_bytes_match_survey():
    assert g.spmv_bytes(1048576, 1048576, 5238784) == 83836932
    assert g.spmm_bytes(1048576, 1048576, 5238784, 64, val=16) == 2256453636
    assert g.spmv_bytes(8000000, 8000000, 213847192) == 2726166308
