"""Boundary behaviour the round-1 review asked for: the single-process multi-GPU handle
(slc_*_upload_sharded / mulv_sharded / mulm_sharded), two host threads sharing one handle, the
accumulating block product, and the FEAST adapter's data flow."""
import threading

import numpy as np
import pytest

import oracle
from helpers import rmat_csc, to_matrix
from sparse_linear_b200 import generators as g
from sparse_linear_b200 import feast, sharded

pytestmark = pytest.mark.gpu


def _ndevs(sl):
    n = sl.device_count()
    return [1] + ([2] if n >= 2 else []) + ([4] if n >= 4 else [])


@pytest.mark.parametrize("dtype", [np.float64, np.complex128])
def test_sharded_handle_products_match_oracle(sl, dtype, monkeypatch):
    monkeypatch.setenv("SLC_SELL_KBASE", "2")
    gm = g.stencil3d_27pt((12, 16, 64), dtype)
    om = oracle.csc(*gm)
    rng = np.random.default_rng(3)
    for ndev in _ndevs(sl):
        sm = sharded.ShardedMatrix(to_matrix(sl, gm), ndev)
        rs, cs = sm.splits()
        assert rs[0] == 0 and rs[-1] == gm.nrows and np.all(np.diff(rs) >= 0) and np.array_equal(rs, cs)
        for step in range(5):                       # both operand buffers, x different every product
            x = rng.uniform(-1, 1, gm.ncols).astype(dtype)
            np.testing.assert_array_equal(sm.mulV(x), oracle.mulv(om, x))
        X = g.uniform_block(gm.ncols, 7, seed=4, dtype=dtype)
        np.testing.assert_array_equal(sm.mulM(X), oracle.mulm(om, X))
        sm.close()
    sl.set_device(0)


def test_sharded_handle_irregular_matrix_allgathers(sl):
    gm = rmat_csc(12, 8, seed=9)
    om = oracle.csc(*gm)
    for ndev in _ndevs(sl):
        sm = sharded.ShardedMatrix(to_matrix(sl, gm), ndev)
        for seed in range(3):
            x = g.uniform_vector(gm.ncols, seed=seed)
            want = oracle.mulv(om, x)
            assert np.linalg.norm(sm.mulV(x) - want) <= 1e-12 * np.linalg.norm(want)
        with pytest.raises(sl.SparseLinearError):
            sm.mulM(g.uniform_block(gm.ncols, 3, seed=1))
        sm.close()
    sl.set_device(0)


def test_second_device_in_the_same_process(sl):
    """Per-device kernel attributes (dynamic shared memory limits) are set per device: everything that
    needed more than 48 KB failed on device 1 in round 1."""
    if sl.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    try:
        sl.set_device(1)
        gm = g.laplace3d_7pt(12)
        m, om = to_matrix(sl, gm), oracle.csc(*gm)
        assert sl.transpose(m) == m                                     # radix scatter: 59 KB dynamic smem
        X = g.uniform_block(gm.ncols, 9, seed=2)
        np.testing.assert_array_equal(sl.mulM(m, X), oracle.mulm(om, X))
        g27 = g.stencil3d_27pt((9, 10, 11))
        m27 = to_matrix(sl, g27)
        np.testing.assert_array_equal(sl.mulM(m27, g.uniform_block(g27.ncols, 9, seed=3)),
                                      oracle.mulm(oracle.csc(*g27), g.uniform_block(g27.ncols, 9, seed=3)))
    finally:
        sl.set_device(0)


def test_two_threads_share_one_handle(sl):
    """Two host threads forcing products on the same handle before any plan exists (what two Haskell
    sparks evaluating `mulV a x` and `mulV a y` do): the handle's mutex serialises plan building and the
    host-operand pipeline."""
    gm = g.stencil3d_27pt((40, 40, 200))      # 320 000 rows: the pipelined host path
    om = oracle.csc(*gm)
    xs = [g.uniform_vector(gm.ncols, seed=s) for s in range(4)]
    wants = [oracle.mulv(om, x) for x in xs]
    Xb = g.uniform_block(gm.ncols, 4, seed=9)
    wantb = oracle.mulm(om, Xb)
    for trial in range(3):
        m = to_matrix(sl, gm)
        m.device()                             # uploaded, unplanned
        errs = []

        def work(i):
            try:
                sl.set_device(0)
                for rep in range(3):
                    if i == 3:
                        if not np.array_equal(sl.mulM(m, Xb), wantb):
                            errs.append(("mulM", rep))
                    elif not np.array_equal(sl.mulV(m, xs[i]), wants[i]):
                        errs.append((i, rep))
            except Exception as e:  # noqa: BLE001
                errs.append(repr(e))
        ts = [threading.Thread(target=work, args=(i,)) for i in range(4)]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        assert not errs, errs


@pytest.mark.parametrize("dtype", [np.float64, np.complex128])
def test_block_product_accumulate_form(sl, dtype):
    """beta = 1 is the reference's axpy_ into a block the caller filled (Sparse.hs:485-488), bit for bit,
    on every SpMM kernel family; other betas start from beta * y."""
    rng = np.random.default_rng(8)
    cx = np.dtype(dtype).kind == "c"
    for gm in (g.laplace2d_5pt(64, dtype), g.stencil3d_27pt((7, 8, 9), dtype), rmat_csc(9, 8, seed=2, dtype=dtype),
               g.laplace2d_5pt(9, dtype)):
        m, om = to_matrix(sl, gm), oracle.csc(*gm)
        k = 6
        X = g.uniform_block(gm.ncols, k, seed=1, dtype=dtype)
        Y0 = g.uniform_block(gm.nrows, k, seed=2, dtype=dtype)
        want = np.stack([oracle.axpy(om, np.ascontiguousarray(X[:, j]), np.ascontiguousarray(Y0[:, j])) for j in range(k)], axis=1)
        got = sl.mulM(m, X, Y0)
        if "sell" in m.device().plan_info():
            np.testing.assert_array_equal(got, want)
        else:
            assert np.linalg.norm(got - want) <= 1e-12 * np.linalg.norm(want)
        beta = (0.5 - 2j) if cx else -3.0
        by = (beta.real * Y0.real - beta.imag * Y0.imag) + 1j * (beta.real * Y0.imag + beta.imag * Y0.real) if cx else beta * Y0
        want = np.stack([oracle.axpy(om, np.ascontiguousarray(X[:, j]), np.ascontiguousarray(by[:, j])) for j in range(k)], axis=1)
        got = sl.mulM(m, X, Y0, beta=beta)
        assert np.linalg.norm(got - want) <= 1e-12 * np.linalg.norm(want)
        np.testing.assert_array_equal(sl.mulM(m, X), oracle.mulm(om, X))      # beta absent: Y is never read
    _ = rng


@pytest.mark.parametrize("dtype", [np.complex128, np.float64])
def test_feast_multiply_work_moving_range(sl, dtype):
    """The RCI data flow of Feast.hs:189-208: host-resident eigenvectors / work1 of m0 columns, ijob 30/40
    with fpm[23], fpm[24] selecting a moving column range, one block product per call; every call is
    compared with the reference loop (oracle.mulm on the active columns), and columns outside the range
    must stay untouched."""
    n_side, m0 = 40, 12
    ga = g.laplace2d_5pt(n_side, dtype)
    gb_host = oracle.ident(ga.ncols, dtype)
    A, B = to_matrix(sl, ga), to_matrix(sl, gb_host)
    oa, ob = oracle.csc(*ga), gb_host
    ws = feast.FeastWorkspace(ga.ncols, m0, dtype)
    ws.eigenvectors[...] = g.uniform_block(ga.ncols, m0, seed=5, dtype=dtype)
    sched = feast.contour_schedule(m0, loops=3) + [(30, m0, 1), (40, 1, 1), (30, 3, 0)]
    for ijob, first, count in sched:
        ws.fpm[23], ws.fpm[24] = first, count
        ws.work1[...] = 7.0                                   # sentinel: inactive columns must keep it
        ws.rci_step(ijob, A, B)
        lo, cnt = first - 1, count
        want = oracle.mulm(oa if ijob == 30 else ob, np.asfortranarray(ws.eigenvectors[:, lo:lo + cnt])) if cnt else None
        if cnt:
            np.testing.assert_array_equal(ws.work1[:, lo:lo + cnt], want)
        assert np.all(ws.work1[:, :lo] == 7.0) and np.all(ws.work1[:, lo + cnt:] == 7.0)
    ws.fpm[23], ws.fpm[24] = m0, 2
    with pytest.raises(sl.SparseLinearError):
        ws.multiplyWork(A)
    ws.close()


@pytest.mark.parametrize("dtype", [np.float64, np.complex128])
def test_device_piece_stacking_and_bucketing(sl, dtype):
    """slc_*_vstack_dev and slc_*_coo_bucket_dev (the device steps of the sharded transpose / assembly)
    against their host restatements, and the single-rank wrappers against the oracle."""
    import ctypes as C
    import torch
    from helpers import random_csc
    rng = np.random.default_rng(31)
    cx = np.dtype(dtype).kind == "c"
    pfx = "z" if cx else "d"
    # vstack: four row blocks of different heights, some empty, one with a long column
    ncols = 57
    blocks = [random_csc(rng, nr, ncols, dens, dtype) for nr, dens in ((40, 0.2), (0, 0.0), (300, 0.4), (17, 0.05))]
    want = sharded.vstack_blocks([g.CSC(*b) for b in blocks])

    def dev(a, dt):
        return torch.from_numpy(np.ascontiguousarray(a).view(dt) if dt is not None else np.ascontiguousarray(a)).cuda()
    got = sharded.vstack_device(pfx, [b.nrows for b in blocks], ncols, [dev(b.colptr.astype(np.int32), None) for b in blocks],
                                [dev(b.rowidx.astype(np.int32), None) for b in blocks],
                                [dev(b.values, np.float64) for b in blocks])
    assert got.nrows == want.nrows and np.array_equal(got.pointers, want.colptr)
    assert np.array_equal(got.indices, want.rowidx) and np.array_equal(got.values, want.values)
    # bucketing: 5 parts, uneven, out-of-range columns go to the nearest part
    n, nc = 5000, 200
    r = rng.integers(0, 90, n).astype(np.int32)
    c = rng.integers(-3, nc + 4, n).astype(np.int32)
    v = rng.uniform(-1, 1, n).astype(dtype)
    if cx:
        v = v + 1j * rng.uniform(-1, 1, n)
    cs = np.array([0, 10, 10, 95, 180, nc], dtype=np.int32)
    host = sharded.compress_pieces(r, c, v, cs)
    dr, dc, dv = dev(r, None), dev(c, None), dev(v, np.float64)
    orow, ocol, oval = torch.empty_like(dr), torch.empty_like(dc), torch.empty_like(dv)
    counts = (C.c_int64 * 5)()
    rc = getattr(sl._capi.lib(), "slc_%s_coo_bucket_dev" % pfx)(
        n, C.c_void_p(dr.data_ptr()), C.c_void_p(dc.data_ptr()), C.c_void_p(dv.data_ptr()), 5, cs.ctypes.data_as(C.c_void_p),
        C.c_void_p(orow.data_ptr()), C.c_void_p(ocol.data_ptr()), C.c_void_p(oval.data_ptr()), counts)
    assert rc == 0, sl._capi.last_error()
    assert [int(x) for x in counts] == [len(h[0]) for h in host]
    vals_back = oval.cpu().numpy().view(dtype) if cx else oval.cpu().numpy()
    at = 0
    for h in host:
        k = len(h[0])
        np.testing.assert_array_equal(orow.cpu().numpy()[at:at + k], h[0])
        np.testing.assert_array_equal(ocol.cpu().numpy()[at:at + k], h[1])
        np.testing.assert_array_equal(vals_back[at:at + k], h[2])
        at += k
    # single-rank wrappers
    m = random_csc(rng, 230, 170, 0.06, dtype)
    t = sharded.sharded_transpose_device(to_matrix(sl, m), np.array([0, m.nrows]), col_splits=np.array([0, m.ncols]))
    ot = oracle.transpose(oracle.csc(*m))
    assert np.array_equal(t.pointers, ot.colptr) and np.array_equal(t.indices, ot.rowidx) and np.array_equal(t.values, ot.values)
    r = rng.integers(0, 230, 3000).astype(np.int32)
    c = rng.integers(0, 170, 3000).astype(np.int32)
    v = rng.integers(-5, 6, 3000).astype(dtype)
    a = sharded.sharded_compress_device(230, dev(r, None), dev(c, None), dev(v, np.float64), np.array([0, 170]), 0, complex_values=cx)
    oa = oracle.compress(230, 170, r, c, v)
    assert np.array_equal(a.pointers, oa.colptr) and np.array_equal(a.indices, oa.rowidx) and np.array_equal(a.values, oa.values)


def test_peer_statistics_reset_clears_waits_sum_and_longest(sl):
    """slc_peer_group_stats(reset=1) must zero all three wait statistics (count, summed ns, longest):
    round 2 cleared the sum, the longest and an unused word, so the count kept the warm-up's waits and
    the mean wait published from it was too small."""
    import ctypes as C
    if sl.device_count() < 2:
        pytest.skip("waits are only recorded against a peer: needs 2 GPUs")
    lib = sl._capi.lib()
    groups = (C.c_void_p * 2)()
    assert lib.slc_peer_group_create_local(2, None, groups) == 0, sl._capi.last_error()
    try:
        for rep in range(5):
            for d in (0, 1):
                sl.set_device(d)
                assert lib.slc_peer_barrier(groups[d]) == 0, sl._capi.last_error()
        for d in (0, 1):
            sl.set_device(d)
            sl.synchronize()
        out = (C.c_int64 * 4)()
        for d in (0, 1):
            assert lib.slc_peer_group_stats(groups[d], out, 1) == 0
            assert out[0] >= 5 and out[1] >= 5 and out[2] > 0 and out[3] > 0, list(out)
            epochs = out[0]
            assert lib.slc_peer_group_stats(groups[d], out, 0) == 0
            assert out[0] == epochs and list(out)[1:] == [0, 0, 0], list(out)   # the epoch count is not a statistic
    finally:
        for d in (0, 1):
            lib.slc_peer_group_free(groups[d])
        sl.set_device(0)
