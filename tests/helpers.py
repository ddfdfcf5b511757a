"""Shared test helpers: the reference's generators restated for numpy, and comparison rules."""
import numpy as np

import oracle


def arbitrary_triples(rng, nr, nc, dtype=np.float64, integer=True):
    """arbitraryMatrix (sparse-linear/tests/Test/LinearAlgebra.hs:29-38): nr*nc/4 + 1 triples with
    uniformly chosen positions, so duplicates occur; values include explicit zeros.  Integer-valued
    by default (the reference runs its properties at Int), so sums are exact in any order."""
    n = nr * nc // 4 + 1
    rows = rng.integers(0, nr, n)
    cols = rng.integers(0, nc, n)
    if integer:
        vals = rng.integers(-9, 10, n).astype(np.float64)
    else:
        vals = rng.uniform(-1, 1, n)
    if np.dtype(dtype).kind == "c":
        im = rng.integers(-9, 10, n).astype(np.float64) if integer else rng.uniform(-1, 1, n)
        vals = vals + 1j * im
    return rows, cols, vals.astype(dtype)


def arbitrary_matrix(rng, nr, nc, dtype=np.float64, integer=True):
    """Oracle-assembled arbitrary matrix as an oracle.CSC."""
    r, c, v = arbitrary_triples(rng, nr, nc, dtype, integer)
    return oracle.compress(nr, nc, r, c, v)


def random_csc(rng, nr, nc, density, dtype=np.float64, integer=False):
    """Random CSC with distinct positions (no duplicates), as an oracle.CSC."""
    nnz = int(nr * nc * density)
    if nnz == 0 or nr == 0 or nc == 0:
        return oracle.csc(nr, nc, np.zeros(nc + 1, np.int32), np.zeros(0, np.int32), np.zeros(0, dtype))
    flat = rng.choice(nr * nc, size=min(nnz, nr * nc), replace=False)
    flat.sort()
    cols, rows = np.divmod(flat, nr)
    colptr = np.zeros(nc + 1, dtype=np.int64)
    np.add.at(colptr, cols + 1, 1)
    colptr = np.cumsum(colptr)
    vals = rng.integers(-9, 10, len(flat)).astype(np.float64) if integer else rng.uniform(-1, 1, len(flat))
    if np.dtype(dtype).kind == "c":
        vals = vals + 1j * (rng.integers(-9, 10, len(flat)) if integer else rng.uniform(-1, 1, len(flat)))
    return oracle.csc(nr, nc, colptr, rows, vals.astype(dtype))


def rmat_csc(scale, edge_factor=8, seed=0, dtype=np.float64):
    """Small power-law matrix: R-MAT triplets assembled by the oracle (duplicates summed)."""
    from sparse_linear_b200 import generators as g
    n, r, c, v = g.rmat_triplets(scale, edge_factor, seed=seed, dtype=dtype)
    return oracle.compress(n, n, r, c, v)


def to_matrix(sl, m):
    """oracle.CSC (or generators.CSC) -> product Matrix"""
    return sl.Matrix(m.ncols, m.nrows, m.colptr, m.rowidx, m.values)


def assert_same_structure(got, want):
    """Bit-exact sparsity structure: the bar for SpGEMM, transpose and assembly."""
    assert got.nrows == want.nrows and got.ncols == want.ncols
    np.testing.assert_array_equal(np.asarray(got.pointers), np.asarray(want.colptr))
    np.testing.assert_array_equal(np.asarray(got.indices), np.asarray(want.rowidx))


def rel_err(got, want):
    got = np.asarray(got).ravel()
    want = np.asarray(want).ravel()
    den = np.linalg.norm(want)
    num = np.linalg.norm(got - want)
    return num / den if den > 0 else num


TOL = 1e-12  # BASELINE.json north_star: values within 1e-12 relative norm-wise error


def assert_values_close(got, want, tol=TOL):
    e = rel_err(got, want)
    assert e <= tol, "relative norm-wise error %.3e > %.1e" % (e, tol)
