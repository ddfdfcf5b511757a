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
    # ALL 64 columns against the reference loop (set 0, axpy_ per column: Feast.hs:203-208), bit for bit
    np.testing.assert_array_equal(Y, oracle.mulm(om, X))
    assert sl.spmm_window_error() == 0


def test_cfg1_and_stencil_complex_spmv(sl):
    """Complex Double mulV at config sizes (the 4-multiply product, Data.Complex), bit-identical."""
    for gm in (g.laplace2d_5pt(1024, np.complex128), g.stencil3d_27pt(100, np.complex128)):
        m, om = to_matrix(sl, gm), oracle.csc(*gm)
        x = g.uniform_vector(gm.ncols, seed=5, dtype=np.complex128)
        np.testing.assert_array_equal(sl.mulV(m, x), oracle.mulv(om, x))
        y0 = g.uniform_vector(gm.nrows, seed=6, dtype=np.complex128)
        np.testing.assert_array_equal(sl.axpy(m, x, y0), oracle.axpy(om, x, y0))


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
    """cfg 4 at scale 20 with the seeded numpy generator (the full scale-24 case is the next test)."""
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


def test_cfg4_rmat_scale24_assembly_spmv_transpose(sl):
    """cfg 4 at its full size: 268,435,456 R-MAT triplets (generated on the GPU: numpy needs minutes),
    assembled on the device and by the oracle -- structure and values bit-identical (duplicates are summed
    in input order by both) --, SpMV within 1e-12 of the oracle (nnz-balanced kernel: rows split across
    lanes), transpose bit-identical."""
    import torch
    import bench
    n, r, c, v = bench.rmat_device(24)
    m = sl.compress_dev(n, n, r.numel(), r.data_ptr(), c.data_ptr(), v.data_ptr(), np.float64)
    rh, ch, vh = r.cpu().numpy(), c.cpu().numpy(), v.cpu().numpy()
    del r, c, v
    torch.cuda.empty_cache()
    om = oracle.compress(n, n, rh, ch, vh)
    del rh, ch, vh
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


def test_cfg5_device_pointer_call_and_spmm_k64(sl):
    """The calls bench.py times, at the benchmarked size: slc_d_mulv_dev (device-resident x, y: one launch
    over all rows, not the chunked host pipeline) and slc_d_mulm_dev with k = 64 on the 200^3 stencil, the
    latter checked on all 64 columns against the reference loop (each oracle column costs 2.4 s; see below
    how 8 of them cover 64)."""
    import ctypes as C
    import torch
    gm = g.stencil3d_27pt(200)
    m, om = to_matrix(sl, gm), oracle.csc(*gm)
    lib = sl._capi.lib()
    x = g.uniform_vector(gm.ncols, seed=3)
    xd = torch.from_numpy(x).cuda()
    yd = torch.full((gm.nrows,), float("nan"), dtype=torch.float64, device="cuda")
    assert lib.slc_d_mulv_dev(m.device().handle, C.c_void_p(xd.data_ptr()), C.c_void_p(yd.data_ptr())) == 0
    sl.synchronize()
    np.testing.assert_array_equal(yd.cpu().numpy(), oracle.mulv(om, x))
    del xd, yd
    # all 64 columns against the reference loop at the price of 8: column j is base vector j % 8 scaled by
    # 2^-(j // 8).  A power-of-two scaling commutes with every rounding of the product (no overflow or
    # subnormals at these magnitudes), so A (2^-s b) == 2^-s (A b) bit for bit and 8 oracle columns give the
    # expected bits of all 64.
    k = 64
    gen = torch.Generator(device="cuda")
    gen.manual_seed(7)
    base = torch.rand(8, gm.ncols, dtype=torch.float64, device="cuda", generator=gen) * 2 - 1
    scale = torch.tensor([2.0 ** -(j // 8) for j in range(k)], dtype=torch.float64, device="cuda")
    X = (base[torch.arange(k, device="cuda") % 8] * scale[:, None]).contiguous()            # column-major n x k
    Y = torch.full((k, gm.nrows), float("nan"), dtype=torch.float64, device="cuda")
    sl.mulM_dev(m, X.data_ptr(), gm.ncols, Y.data_ptr(), gm.nrows, k)
    sl.synchronize()
    want8 = oracle.mulm(om, np.asfortranarray(base.cpu().numpy().T))                        # n x 8
    got = Y.cpu().numpy()                                                                    # k x n
    for j in range(k):
        np.testing.assert_array_equal(got[j], want8[:, j % 8] * 2.0 ** -(j // 8), err_msg="column %d" % j)
    assert not bool(torch.isnan(Y).any())


def test_stencil27_64_spmm_all_64_columns(sl):
    gm = g.stencil3d_27pt(64)
    m, om = to_matrix(sl, gm), oracle.csc(*gm)
    X = g.uniform_block(gm.ncols, 64, seed=11)
    np.testing.assert_array_equal(sl.mulM(m, X), oracle.mulm(om, X))
