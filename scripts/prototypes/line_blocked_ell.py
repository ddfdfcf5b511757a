"""Design study for the next SpMM kernel ("Next" item 2): line-blocked ELL.

k_spmm_panel is bound by the L1 tag stage: every (row, column-of-X) product costs one 8-byte gather.  On
a stencil the rows r and r + S (S = distance between grid lines) share most of their columns, so a thread
that owns R such rows can load each distinct column once and feed up to R accumulators from it.  This
module builds that format on the host with numpy and evaluates the product exactly as a kernel would
(one pass over the thread's merged column list, multiply and add rounded separately, every row's terms in
increasing column order), so that tests can check two things before any CUDA is written:

  * the result is bit-identical to the reference order (oracle.mulm), for any matrix, any S and R;
  * the load count per product on the stencils of BASELINE.json (27-point: 36 loads for 54 products at
    R = 2, 54 for 108 at R = 4).

Nothing here is on the product path; it is numpy on small matrices.
"""
import numpy as np


def build(rowptr, colidx, vals, nrows, stride, rows_per_thread):
    """Row groups {p + i * stride : i < R} for p in every block of R lines; returns a list of
    (rows, cols, mask, coef): rows[R] (-1 = none), cols[U] merged ascending, mask[U, R] bool, coef[U, R]."""
    R = rows_per_thread
    groups = []
    block = R * stride
    for b0 in range(0, nrows, block):
        for p in range(b0, min(b0 + stride, nrows)):
            rows = [p + i * stride if p + i * stride < min(nrows, b0 + block) else -1 for i in range(R)]
            lists = [colidx[rowptr[r]:rowptr[r + 1]] if r >= 0 else np.zeros(0, colidx.dtype) for r in rows]
            cols = np.unique(np.concatenate(lists)) if any(len(x) for x in lists) else np.zeros(0, colidx.dtype)
            mask = np.zeros((len(cols), R), dtype=bool)
            coef = np.zeros((len(cols), R), dtype=vals.dtype)
            for i, r in enumerate(rows):
                if r < 0:
                    continue
                at = np.searchsorted(cols, lists[i])
                mask[at, i] = True
                coef[at, i] = vals[rowptr[r]:rowptr[r + 1]]
            groups.append((rows, cols, mask, coef))
    return groups


def multiply(groups, nrows, X):
    """Y = A X evaluated group by group: one load of X[c, :] per merged column, then the accumulators of the
    rows that hold c; a * x and acc + (a * x) are separate IEEE operations, as in the reference."""
    Y = np.zeros((nrows, X.shape[1]), dtype=X.dtype, order="F")
    loads = products = 0
    for rows, cols, mask, coef in groups:
        acc = np.zeros((len(rows), X.shape[1]), dtype=X.dtype)
        for u, c in enumerate(cols):
            x = X[c, :]                      # the one gather
            loads += 1
            for i in range(len(rows)):
                if mask[u, i]:
                    acc[i] = coef[u, i] * x + acc[i]
                    products += 1
        for i, r in enumerate(rows):
            if r >= 0:
                Y[r] = acc[i]
    return Y, loads, products
