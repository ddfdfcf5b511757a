"""Seeded synthetic inputs for BASELINE.json's configs (host side, numpy only).

Matrices come back as plain CSC arrays (nrows, ncols, colptr, rowidx, values) built directly in
O(nnz); nothing here goes through the reference's quadratic `glin`.  The kronecker-sum
construction the configs name (Data/Matrix/Sparse.hs:597-638 + :401-431) gives exactly these
matrices; tests/test_generators.py checks that on small grids.
"""
from collections import namedtuple

import numpy as np

CSC = namedtuple("CSC", "nrows ncols colptr rowidx values")


def _stencil(shape, offsets, weights, dtype, col_range=None):
    """Structurally symmetric stencil on a row-major grid.  offsets: list of index tuples.
    col_range=(c0, c1) builds only the column block A[:, c0:c1] (n rows, c1-c0 columns)."""
    shape = tuple(int(s) for s in shape)
    nd = len(shape)
    ntot = int(np.prod(shape))
    c0, c1 = (0, ntot) if col_range is None else (int(col_range[0]), int(col_range[1]))
    n = c1 - c0
    strides = [int(np.prod(shape[d + 1:])) for d in range(nd)]
    lin = [sum(o[d] * strides[d] for d in range(nd)) for o in offsets]
    order = np.argsort(lin, kind="stable")
    offsets = [offsets[i] for i in order]
    weights = [weights[i] for i in order]
    lin = [lin[i] for i in order]
    if len(set(lin)) != len(lin):
        raise ValueError("grid too small for this stencil: offsets alias")
    coords = np.unravel_index(np.arange(c0, c1, dtype=np.int64), shape)
    k = len(offsets)
    valid = np.empty((n, k), dtype=bool)
    for j, o in enumerate(offsets):
        ok = np.ones(n, dtype=bool)
        for d in range(nd):
            if o[d]:
                cd = coords[d] + o[d]
                ok &= (cd >= 0) & (cd < shape[d])
        valid[:, j] = ok
    del coords
    counts = valid.sum(axis=1, dtype=np.int64)
    colptr = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(counts, out=colptr[1:])
    # entry (c, j): row = c + lin[j]   (column c; symmetric offset set => same as row view)
    rows = (np.arange(c0, c1, dtype=np.int32)[:, None] + np.asarray(lin, dtype=np.int32)[None, :])[valid]
    vals = np.broadcast_to(np.asarray(weights, dtype=dtype)[None, :], (n, k))[valid]
    return CSC(ntot, n, colptr.astype(np.int32), np.ascontiguousarray(rows, dtype=np.int32),
               np.ascontiguousarray(vals, dtype=dtype))


def laplace2d_5pt(n, dtype=np.float64):
    """cfg 1/2: A = kron(I,T) + kron(T,I), T = tridiag(-1, 2, -1) of order n."""
    offs = [(0, 0), (0, 1), (0, -1), (1, 0), (-1, 0)]
    w = [4.0, -1.0, -1.0, -1.0, -1.0]
    return _stencil((n, n), offs, w, dtype)


def laplace3d_7pt(n, dtype=np.float64):
    """cfg 3: kron(kron(I,I),T) + kron(kron(I,T),I) + kron(kron(T,I),I)."""
    offs = [(0, 0, 0), (0, 0, 1), (0, 0, -1), (0, 1, 0), (0, -1, 0), (1, 0, 0), (-1, 0, 0)]
    w = [6.0] + [-1.0] * 6
    return _stencil((n, n, n), offs, w, dtype)


def stencil3d_27pt(shape, dtype=np.float64, col_range=None):
    """cfg 5: 27-point stencil, centre 26, off-centre -1.  shape = n or (nz, ny, nx)."""
    if np.isscalar(shape):
        shape = (int(shape),) * 3
    offs, w = [], []
    for dz in (-1, 0, 1):
        for dy in (-1, 0, 1):
            for dx in (-1, 0, 1):
                offs.append((dz, dy, dx))
                w.append(26.0 if (dz, dy, dx) == (0, 0, 0) else -1.0)
    return _stencil(shape, offs, w, dtype, col_range)


def tridiag(n, lo=-1.0, d=2.0, up=-1.0, dtype=np.float64):
    return _stencil((n,), [(0,), (1,), (-1,)], [d, lo, up], dtype)


def rmat_triplets(scale, edge_factor=16, abcd=(0.57, 0.19, 0.19, 0.05), seed=0, dtype=np.float64):
    """cfg 4: R-MAT edge list (rows, cols, vals) with duplicates, values uniform(-1, 1)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    n = 1 << scale
    m = n * edge_factor
    a, b, c, _ = abcd
    rows = np.zeros(m, dtype=np.int32)
    cols = np.zeros(m, dtype=np.int32)
    for bit in range(scale):
        u = rng.random(m, dtype=np.float32)
        rbit = u >= (a + b)                       # quadrants c, d: lower half
        cbit = ((u >= a) & (u < a + b)) | (u >= a + b + c)  # quadrants b, d: right half
        rows |= rbit.astype(np.int32) << bit
        cols |= cbit.astype(np.int32) << bit
    vals = rng.uniform(-1.0, 1.0, m).astype(dtype)
    if np.dtype(dtype).kind == "c":
        vals = vals + 1j * rng.uniform(-1.0, 1.0, m)
    return n, rows, cols, vals


def uniform_vector(n, seed=0, dtype=np.float64):
    rng = np.random.Generator(np.random.PCG64(seed))
    x = rng.uniform(-1.0, 1.0, n)
    if np.dtype(dtype).kind == "c":
        x = x + 1j * rng.uniform(-1.0, 1.0, n)
    return np.ascontiguousarray(x, dtype=dtype)


def uniform_block(n, k, seed=0, dtype=np.float64):
    """Dense (n, k) block, column-major, re and im uniform(-1, 1)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    X = np.empty((n, k), dtype=dtype, order="F")
    for j in range(k):
        col = rng.uniform(-1.0, 1.0, n)
        if np.dtype(dtype).kind == "c":
            col = col + 1j * rng.uniform(-1.0, 1.0, n)
        X[:, j] = col
    return X


def split_duplicates(m, seed=0):
    """cfg 3 assembly input: every entry v of m becomes 0.25 v and 0.75 v (which sum back to v
    exactly), concatenated and shuffled.  Returns (rows, cols, vals)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    cols = np.repeat(np.arange(m.ncols, dtype=np.int32), np.diff(m.colptr))
    rows2 = np.concatenate([m.rowidx, m.rowidx])
    cols2 = np.concatenate([cols, cols])
    vals2 = np.concatenate([0.25 * m.values, 0.75 * m.values])
    perm = rng.permutation(len(rows2))
    return rows2[perm], cols2[perm], vals2[perm]


# algorithmic bytes; idx = ptr = 4 B, val = 8 or 16 B
def spmv_bytes(nrows, ncols, nnz, val=8):
    return nnz * (val + 4) + (nrows + 1) * 4 + ncols * val + nrows * val


def spmm_bytes(nrows, ncols, nnz, k, val=8):
    return nnz * (val + 4) + (nrows + 1) * 4 + ncols * k * val + nrows * k * val
