"""The oracle against everything the reference's own tests pin for the hot path (CPU only).

Known answers: sparse-linear/tests/Sparse.hs:61-73 (tests/golden/reference_known_answers.json).
Properties: Sparse.hs:23-165 and Test/LinearAlgebra.hs:29-67, re-run on the oracle with the same
generator shape.  scipy is an independent second opinion, never the definition of truth.
"""
import json
import os

import numpy as np
import pytest
import scipy.sparse as sp

import oracle
from helpers import arbitrary_matrix, arbitrary_triples, random_csc

GOLD = os.path.join(os.path.dirname(__file__), "golden")
DTYPES = [np.float64, np.complex128]


def same(a, b):
    return (a.nrows == b.nrows and a.ncols == b.ncols and np.array_equal(a.colptr, b.colptr)
            and np.array_equal(a.rowidx, b.rowidx) and bool(np.all(a.values == b.values)))


def to_scipy(m):
    return sp.csc_matrix((m.values, m.rowidx, m.colptr), shape=(m.nrows, m.ncols))


def dims(rng, lo=1, hi=12):
    return int(rng.integers(lo, hi + 1))


# ---- known answers from the reference's test source ---------------------------------------------

def test_reference_ctrans_known_answers():
    cases = json.load(open(os.path.join(GOLD, "reference_known_answers.json")))["ctrans_fixed_points"]
    for c in cases:
        t = c["triples"]
        m = oracle.from_triples(c["nrows"], c["ncols"], [(r, col, complex(re, im)) for r, col, re, im in t])
        assert same(oracle.ctrans(m), m), c["name"]
        assert oracle.hermitian(m), c["name"]


def test_reference_feast_matrix_eigenvalues():
    c = json.load(open(os.path.join(GOLD, "reference_known_answers.json")))["feast_matrix"]
    m = oracle.from_triples(2, 2, [(r, col, complex(re, im)) for r, col, re, im in c["triples"]])
    assert oracle.hermitian(m)
    dense = np.column_stack([oracle.mulv(m, e) for e in np.eye(2, dtype=np.complex128)])
    ev = np.linalg.eigvalsh(dense)
    assert np.allclose(ev, c["eigenvalues"], atol=c["tolerance"])


# ---- properties of the reference's test-suite ------------------------------------------------------

@pytest.mark.parametrize("seed", range(25))
def test_from_triples_format_properties(seed):  # Sparse.hs:23, LinearAlgebra.hs:40-67
    rng = np.random.default_rng(seed)
    m = arbitrary_matrix(rng, dims(rng), dims(rng))
    assert oracle.check_format(m) == 0


@pytest.mark.parametrize("seed", range(10))
def test_mulv_identity(seed):  # Sparse.hs:41-47
    rng = np.random.default_rng(seed)
    n = dims(rng, 1, 40)
    v = rng.integers(-50, 50, n).astype(np.float64)
    assert np.array_equal(oracle.mulv(oracle.ident(n), v), v)


@pytest.mark.parametrize("seed", range(15))
def test_mm_identities_and_assoc(seed):  # Sparse.hs:75-102
    rng = np.random.default_rng(seed)
    a = arbitrary_matrix(rng, dims(rng), dims(rng))
    assert same(oracle.mm(oracle.ident(a.nrows), a), a)
    assert same(oracle.mm(a, oracle.ident(a.ncols)), a)
    n, p = dims(rng), dims(rng)
    b = arbitrary_matrix(rng, a.ncols, n)
    c = arbitrary_matrix(rng, n, p)
    left = oracle.mm(oracle.mm(a, b), c)
    right = oracle.mm(a, oracle.mm(b, c))
    assert same(left, right)
    assert oracle.check_format(left) == 0


@pytest.mark.parametrize("seed", range(15))
def test_addition_properties(seed):  # Sparse.hs:147-165
    rng = np.random.default_rng(seed)
    a = arbitrary_matrix(rng, dims(rng), dims(rng))
    b = arbitrary_matrix(rng, a.nrows, a.ncols)
    c = arbitrary_matrix(rng, a.nrows, a.ncols)
    z = oracle.csc(a.nrows, a.ncols, np.zeros(a.ncols + 1), [], np.zeros(0))
    assert same(oracle.add(a, z), a)                                   # a + zeros == a
    amina = oracle.lin(1, a, -1, a)
    assert same(amina, oracle.csc(a.nrows, a.ncols, a.colptr, a.rowidx, a.values * 0))  # a - a == 0*a
    assert same(oracle.add(a, b), oracle.add(b, a))
    assert same(oracle.add(a, oracle.add(b, c)), oracle.add(oracle.add(a, b), c))
    assert oracle.check_format(oracle.add(a, b)) == 0


@pytest.mark.parametrize("seed", range(8))
def test_transpose_diag_and_kronecker_ident(seed):  # Sparse.hs:26-29,56-59
    rng = np.random.default_rng(seed)
    n, k = dims(rng), dims(rng)
    d = oracle.diag(rng.integers(-9, 10, n).astype(np.float64))
    assert same(oracle.transpose(d), d)
    assert same(oracle.kronecker(oracle.ident(n), oracle.ident(k)), oracle.ident(n * k))
    a, b = arbitrary_matrix(rng, dims(rng, 1, 6), dims(rng, 1, 6)), arbitrary_matrix(rng, dims(rng, 1, 6), dims(rng, 1, 6))
    assert oracle.check_format(oracle.kronecker(a, b)) == 0


@pytest.mark.parametrize("seed", range(8))
def test_symmetric_blockwise_ctrans(seed):  # Sparse.hs:115-127, restated without fromBlocksDiag
    rng = np.random.default_rng(seed)
    n = dims(rng)
    m = arbitrary_matrix(rng, n, n, np.complex128)
    s = oracle.add(m, oracle.ctrans(m))
    assert same(oracle.ctrans(s), s)
    assert oracle.hermitian(s)


# ---- the two accumulators agree bit for bit ---------------------------------------------------------

@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("seed", range(6))
def test_literal_and_fast_accumulators_agree(seed, dtype):
    rng = np.random.default_rng(100 + seed)
    a = arbitrary_matrix(rng, dims(rng, 1, 30), dims(rng, 1, 30), dtype, integer=False)
    b = arbitrary_matrix(rng, a.ncols, dims(rng, 1, 30), dtype, integer=False)
    assert same(oracle.mm(a, b), oracle.mm(a, b, literal=True))
    c = arbitrary_matrix(rng, a.nrows, a.ncols, dtype, integer=False)
    assert same(oracle.lin(0.5, a, -2.0, c), oracle.lin(0.5, a, -2.0, c, literal=True))


# ---- independent cross-check against scipy ------------------------------------------------------------

@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("seed", range(5))
def test_against_scipy(seed, dtype):
    rng = np.random.default_rng(200 + seed)
    a = random_csc(rng, 37, 29, 0.15, dtype)
    b = random_csc(rng, 29, 23, 0.2, dtype)
    A, B = to_scipy(a), to_scipy(b)
    x = rng.uniform(-1, 1, 29).astype(dtype)
    assert np.allclose(oracle.mulv(a, x), A @ x, rtol=1e-13, atol=1e-13)
    X = rng.uniform(-1, 1, (29, 5)).astype(dtype)
    assert np.allclose(oracle.mulm(a, X), A @ X, rtol=1e-13, atol=1e-13)
    t = oracle.transpose(a)
    T = A.T.tocsc()
    T.sort_indices()
    assert np.array_equal(t.colptr, T.indptr) and np.array_equal(t.rowidx, T.indices)
    assert np.array_equal(t.values, T.data)
    c = oracle.mm(a, b)
    Cs = (A @ B).tocsc()
    Cs.sort_indices()
    # scipy may drop nothing here (random values never cancel exactly), so structures coincide
    assert np.array_equal(c.colptr, Cs.indptr) and np.array_equal(c.rowidx, Cs.indices)
    assert np.allclose(c.values, Cs.data, rtol=1e-13, atol=1e-13)
    r, cc, v = arbitrary_triples(rng, 17, 13, dtype, integer=False)
    m = oracle.compress(17, 13, r, cc, v)
    S = sp.coo_matrix((v, (r, cc)), shape=(17, 13)).tocsc()
    S.sum_duplicates()
    S.sort_indices()
    assert np.array_equal(m.colptr, S.indptr) and np.array_equal(m.rowidx, S.indices)
    assert np.allclose(m.values, S.data, rtol=1e-13, atol=1e-13)


# ---- semantics the GPU path must reproduce --------------------------------------------------------------

def test_structural_product_keeps_cancellations_and_explicit_zeros():
    # [1 1] * [1; -1] = [0]: the entry exists although it is numerically zero
    a = oracle.from_triples(1, 2, [(0, 0, 1.0), (0, 1, 1.0)])
    b = oracle.from_triples(2, 1, [(0, 0, 1.0), (1, 0, -1.0)])
    c = oracle.mm(a, b)
    assert c.colptr.tolist() == [0, 1] and c.rowidx.tolist() == [0] and c.values.tolist() == [0.0]
    # explicit zero in an operand still creates structure
    z = oracle.from_triples(2, 2, [(0, 0, 0.0), (1, 1, 5.0)])
    c2 = oracle.mm(z, z)
    assert c2.colptr.tolist() == [0, 1, 2]


def test_compress_keeps_zeros_sums_in_input_order_and_reports_bounds():
    m = oracle.compress(3, 3, [2, 0, 2, 2], [1, 0, 1, 1], [1e16, 0.0, 1.0, -1e16])
    assert m.colptr.tolist() == [0, 1, 2, 2]
    assert m.rowidx.tolist() == [0, 2]
    assert m.values.tolist() == [0.0, (1e16 + 1.0) + -1e16]   # input order: ((1e16 + 1) - 1e16)
    with pytest.raises(oracle.OracleError) as e:
        oracle.compress(3, 3, [0, 3, 5], [0, 9, 0], [1.0, 1.0, 1.0])
    assert e.value.code == oracle.ERR_ROW_BOUNDS and e.value.pos == 1   # rows are checked first
    with pytest.raises(oracle.OracleError) as e:
        oracle.compress(3, 3, [0, 1, 2], [0, 9, -1], [1.0, 1.0, 1.0])
    assert e.value.code == oracle.ERR_COL_BOUNDS and e.value.pos == 1
    with pytest.raises(oracle.OracleError):
        oracle.compress(3, 3, [0, 1], [0], [1.0, 1.0])


def test_spmv_order_is_column_order_no_fma():
    # row 0 = [1e16, 1, -1e16]: column order gives ((1e16*1 + 0) + 1) - 1e16 = 0, not 1
    m = oracle.from_triples(1, 3, [(0, 0, 1e16), (0, 1, 1.0), (0, 2, -1e16)])
    assert oracle.mulv(m, np.ones(3)).tolist() == [0.0]
    # a*x + y with separately rounded product: (1+2^-30)^2 rounds away the 2^-60 term
    a = 1.0 + 2.0 ** -30
    m2 = oracle.from_triples(1, 1, [(0, 0, a)])
    got = oracle.axpy(m2, np.array([a]), np.array([-(1.0 + 2.0 ** -29)]))
    assert got.tolist() == [0.0]   # an FMA would give 2^-60


def test_mm_dimension_error():
    a = oracle.ident(3)
    b = oracle.ident(4)
    with pytest.raises(oracle.OracleError, match="mm: inner dimension mismatch"):
        oracle.mm(a, b)
    with pytest.raises(oracle.OracleError, match="glin: row number mismatch"):
        oracle.add(a, b)


def test_decompress():
    assert oracle.decompress([0, 2, 2, 5]).tolist() == [0, 0, 2, 2, 2]
