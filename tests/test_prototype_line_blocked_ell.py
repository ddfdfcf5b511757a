"""The line-blocked ELL design study (scripts/prototypes/line_blocked_ell.py): merged column lists keep the
reference's per-row summation order bit for bit, and cut the loads per product as DESIGN.md says."""
import os
import sys

import numpy as np
import pytest

import oracle
from helpers import random_csc
from sparse_linear_b200 import generators as g

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scripts", "prototypes"))
import line_blocked_ell as lb  # noqa: E402


def _rows(m):
    t = oracle.transpose(oracle.csc(*m))      # CSC of the transpose = rows of m
    return np.asarray(t.colptr, dtype=np.int64), np.asarray(t.rowidx), np.asarray(t.values)


@pytest.mark.parametrize("dtype", [np.float64, np.complex128])
def test_any_matrix_any_stride_is_bit_exact(dtype):
    rng = np.random.default_rng(3)
    for m, stride, R in ((random_csc(rng, 37, 41, 0.2, dtype), 5, 2), (random_csc(rng, 64, 64, 0.1, dtype), 8, 4),
                         (g.laplace2d_5pt(9, dtype), 9, 3), (random_csc(rng, 10, 7, 0.0, dtype), 3, 2)):
        rp, ci, va = _rows(m)
        X = rng.uniform(-1, 1, (m.ncols, 3)).astype(dtype)
        if np.dtype(dtype).kind == "c":
            X = X + 1j * rng.uniform(-1, 1, (m.ncols, 3))
        X = np.asfortranarray(X)
        Y, _, products = lb.multiply(lb.build(rp, ci, va, m.nrows, stride, R), m.nrows, X)
        assert products == int(rp[-1])
        if np.dtype(dtype).kind == "c":
            # numpy's complex multiply may fuse differently from the 4-multiply formula: compare to tolerance
            want = oracle.mulm(oracle.csc(*m), X)
            assert np.linalg.norm(Y - want) <= 1e-14 * max(np.linalg.norm(want), 1e-300)
        else:
            np.testing.assert_array_equal(Y, oracle.mulm(oracle.csc(*m), X))


def test_load_reduction_on_the_27_point_stencil():
    shape = (6, 7, 16)                        # lines of 16 rows
    m = g.stencil3d_27pt(shape)
    rp, ci, va = _rows(m)
    X = np.asfortranarray(np.random.default_rng(0).uniform(-1, 1, (m.ncols, 2)))
    ratios = {}
    for R in (1, 2, 4):
        Y, loads, products = lb.multiply(lb.build(rp, ci, va, m.nrows, 16, R), m.nrows, X)
        np.testing.assert_array_equal(Y, oracle.mulm(oracle.csc(*m), X))
        ratios[R] = loads / products
    assert ratios[1] == 1.0
    # interior: 36 loads for 54 products at R = 2, 54 for 108 at R = 4; grid boundaries cost a little more
    assert 0.66 <= ratios[2] <= 0.72
    assert 0.50 <= ratios[4] <= 0.60
