"""The reference's QuickCheck properties (sparse-linear/tests/Sparse.hs:23-165) run through the
GPU path, with the reference's generator shape (Test/LinearAlgebra.hs:29-38) and integer-valued
doubles where the reference used Int, so equality is exact."""
import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

import oracle
from helpers import arbitrary_triples

pytestmark = pytest.mark.gpu
SETTINGS = dict(max_examples=25, deadline=None, suppress_health_check=list(HealthCheck))
dim = st.integers(min_value=1, max_value=14)
seed = st.integers(min_value=0, max_value=2 ** 31 - 1)


def arb(sl, s, nr, nc, dtype=np.float64):
    r, c, v = arbitrary_triples(np.random.default_rng(s), nr, nc, dtype)
    return sl.compress(nr, nc, r, c, v)


def check_matrix(m):
    """checkMatrix, Test/LinearAlgebra.hs:40-67"""
    assert oracle.check_format(oracle.csc(m.nrows, m.ncols, m.pointers, m.indices, m.values)) == 0


@settings(**SETTINGS)
@given(seed, dim, dim)
def test_from_triples_format(sl, s, nr, nc):
    check_matrix(arb(sl, s, nr, nc))


@settings(**SETTINGS)
@given(dim, dim)
def test_kronecker_identities(sl, m, n):
    assert sl.kronecker(sl.ident(m), sl.ident(n)) == sl.ident(m * n)


@settings(**SETTINGS)
@given(seed, dim, dim, dim, dim)
def test_kronecker_format(sl, s, a, b, c, d):
    check_matrix(sl.kronecker(arb(sl, s, a, b), arb(sl, s + 1, c, d)))


@settings(**SETTINGS)
@given(seed, st.integers(1, 200))
def test_mulv_identity(sl, s, m):
    v = np.random.default_rng(s).integers(-1000, 1000, m).astype(np.float64)
    assert np.array_equal(sl.mulV(sl.ident(m), v), v)


@settings(**SETTINGS)
@given(seed, dim, dim)
def test_addition_laws(sl, s, nr, nc):
    a, b, c = arb(sl, s, nr, nc), arb(sl, s + 1, nr, nc), arb(sl, s + 2, nr, nc)
    assert a + sl.zeros(nr, nc) == a
    assert a - a == sl.cmap(lambda v: v * 0, a)
    assert a + b == b + a
    assert a + (b + c) == (a + b) + c
    check_matrix(a + b)


@settings(**SETTINGS)
@given(seed, st.integers(1, 60))
def test_transpose_diag(sl, s, n):
    d = sl.diag(np.random.default_rng(s).integers(-9, 10, n).astype(np.float64))
    assert sl.transpose(d) == d


@settings(**SETTINGS)
@given(seed, dim, dim, dim, dim)
def test_mul_laws(sl, s, m, n, p, q):
    a = arb(sl, s, m, n)
    assert sl.ident(m) * a == a
    assert a * sl.ident(n) == a
    b, c = arb(sl, s + 1, n, p), arb(sl, s + 2, p, q)
    assert (a * b) * c == a * (b * c)
    check_matrix(a * b)


@settings(**SETTINGS)
@given(seed, dim)
def test_symmetric_via_ctrans(sl, s, n):
    m = arb(sl, s, n, n, np.complex128)
    sym = m + sl.ctrans(m)
    assert sl.ctrans(sym) == sym
    assert sl.hermitian(sym)
