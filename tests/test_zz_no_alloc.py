"""slc_coo_bucket_noalloc_dev: the triplet bucketing for hosts that may not allocate device memory.

Some users of this library cannot let a call reach cudaMalloc: a solver that has already given all of
HBM to its own arena, a Haskell `unsafe` foreign call or a service loop that cannot wait on a driver
allocation, a GPU shared with a runtime whose allocator keeps every free byte.  For them the library
carries a fixed work area inside its own image (a __device__ array, about 0.5 MiB per GPU) and this entry
point buckets triplets in it instead of in pool memory.  The price is a size limit
(SLC_NOALLOC_MAX_TRIPLETS per call, 16 parts) and one shared area per GPU, which still holds the previous
call's numbers when the next call starts.

The checks here go through the C ABI only.  The allocating entry points slc_{d,z}_coo_bucket_dev are the
yardstick: rows, rebased columns, values and counts must be identical, bit for bit (the work is integer
moves, so the tolerance is zero, as in the other bucketing test).  Every case makes several calls in a row
with different sizes and part tables, so a call that trusted leftovers in the work area would be caught.

This file sorts last on purpose (pytest -x stops at the first failure; nothing should hide behind it) and
needs nothing from the tests before it: it passes alone in a fresh process and after the whole suite.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_TRIPLETS = 1 << 24          # SLC_NOALLOC_MAX_TRIPLETS in include/slc.h
ERR_RANGE = -8


def _dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _triplets(rng, n, nrows, ncols, dtype, stray=3):
    """n triplets; a few columns lie outside [0, ncols) on either side (they go to the nearest part)."""
    r = rng.integers(0, nrows, n).astype(np.int32)
    c = rng.integers(-stray, ncols + stray + 1, n).astype(np.int32)
    v = rng.uniform(-1, 1, n)
    if np.dtype(dtype).kind == "c":
        v = v + 1j * rng.uniform(-1, 1, n)
    return r, c, v.astype(dtype)


def _bucket(sl, which, r, c, v, splits):
    """One call through the C ABI.  which = 'noalloc' or 'alloc'.  Returns host copies of
    (rows, cols, vals as float64 words, counts)."""
    import torch
    cx = v.dtype.kind == "c"
    n, nparts = len(r), len(splits) - 1
    dr, dc, dv = _dev(r), _dev(c), _dev(v.view(np.float64))
    # outputs start from a sentinel so that a position nobody wrote shows up
    orow = torch.full_like(dr, -7)
    ocol = torch.full_like(dc, -7)
    oval = torch.full_like(dv, float("nan"))
    counts = (C.c_int64 * nparts)(*([-1] * nparts))
    cs = np.ascontiguousarray(splits, dtype=np.int32)
    ptrs = [C.c_void_p(t.data_ptr()) for t in (dr, dc, dv)]
    outs = [C.c_void_p(t.data_ptr()) for t in (orow, ocol, oval)]
    lib = sl._capi.lib()
    if which == "noalloc":
        rc = lib.slc_coo_bucket_noalloc_dev(ord("z" if cx else "d"), n, *ptrs, nparts, cs.ctypes.data_as(C.c_void_p),
                                            *outs, counts)
    else:
        rc = getattr(lib, "slc_%s_coo_bucket_dev" % ("z" if cx else "d"))(
            n, *ptrs, nparts, cs.ctypes.data_as(C.c_void_p), *outs, counts)
    assert rc == 0, sl._capi.last_error()
    torch.cuda.synchronize()
    return orow.cpu().numpy(), ocol.cpu().numpy(), oval.cpu().numpy(), [int(x) for x in counts]


def _same(a, b):
    assert a[3] == b[3]
    np.testing.assert_array_equal(a[0], b[0])
    np.testing.assert_array_equal(a[1], b[1])
    # compare the bit patterns: the values are moved, never computed with
    np.testing.assert_array_equal(a[2].view(np.uint64), b[2].view(np.uint64))


# (triplets, rows, columns, column bounds of the parts): consecutive calls differ in size and in the
# number of parts, the second is smaller than the first (stale tile counters beyond it), the third has
# an empty part and a single-column part, the fourth is one part only, the last repeats the first
CASES = [
    (300_001, 5000, 4096, [0, 512, 1024, 1536, 2048, 2560, 3072, 3584, 4096]),
    (4_097, 90, 200, [0, 10, 10, 95, 180, 200]),
    (70_000, 300, 64, [0, 1, 1, 2, 33, 64]),
    (2_048, 10, 10, [0, 10]),
    (300_001, 5000, 4096, [0, 512, 1024, 1536, 2048, 2560, 3072, 3584, 4096]),
]


@pytest.mark.parametrize("dtype", [np.float64, np.complex128])
def test_noalloc_bucketing_matches_allocating_entry_point_call_after_call(sl, dtype):
    rng = np.random.default_rng(2026)
    for n, nrows, ncols, splits in CASES:
        r, c, v = _triplets(rng, n, nrows, ncols, dtype)
        want = _bucket(sl, "alloc", r, c, v, splits)
        got1 = _bucket(sl, "noalloc", r, c, v, splits)
        got2 = _bucket(sl, "noalloc", r, c, v, splits)     # straight after: the work area is dirty
        _same(got1, want)
        _same(got2, want)
        assert sum(want[3]) == n


def test_noalloc_bucketing_sixteen_parts_and_input_order(sl):
    """The widest part table the work area holds, and the contract spelled out independently of the
    other entry point: inside a part the triplets keep their input order, columns are rebased."""
    rng = np.random.default_rng(7)
    n, ncols = 123_457, 1600
    splits = np.arange(0, ncols + 1, 100, dtype=np.int32)          # 16 parts
    r, c, v = _triplets(rng, n, 977, ncols, np.float64, stray=0)
    c = np.minimum(c, ncols - 1).astype(np.int32)
    rows, cols, vals, counts = _bucket(sl, "noalloc", r, c, v, splits)
    owner = c // 100
    order = np.argsort(owner, kind="stable")
    np.testing.assert_array_equal(rows, r[order])
    np.testing.assert_array_equal(cols, (c - owner * 100)[order])
    np.testing.assert_array_equal(vals, v[order])
    assert counts == np.bincount(owner, minlength=16).tolist()


def test_noalloc_bucketing_limits_and_empty_input(sl):
    import torch
    lib = sl._capi.lib()
    one = torch.zeros(4, dtype=torch.int32, device="cuda")
    val = torch.zeros(4, dtype=torch.float64, device="cuda")
    p = C.c_void_p(one.data_ptr())
    pv = C.c_void_p(val.data_ptr())
    cs = np.array([0, 5, 9], dtype=np.int32)
    counts = (C.c_int64 * 17)()
    # one triplet too many: refused before anything is read or written
    rc = lib.slc_coo_bucket_noalloc_dev(ord("d"), MAX_TRIPLETS + 1, p, p, pv, 2, cs.ctypes.data_as(C.c_void_p), p, p, pv, counts)
    assert rc == ERR_RANGE
    assert str(MAX_TRIPLETS) in sl._capi.last_error()
    # 17 parts
    cs17 = np.arange(18, dtype=np.int32)
    rc = lib.slc_coo_bucket_noalloc_dev(ord("d"), 4, p, p, pv, 17, cs17.ctypes.data_as(C.c_void_p), p, p, pv, counts)
    assert rc == ERR_RANGE and "16" in sl._capi.last_error()
    # unknown element type
    rc = lib.slc_coo_bucket_noalloc_dev(ord("s"), 4, p, p, pv, 2, cs.ctypes.data_as(C.c_void_p), p, p, pv, counts)
    assert rc == -1
    # nothing to do
    counts[0], counts[1] = 5, 5
    rc = lib.slc_coo_bucket_noalloc_dev(ord("z"), 0, None, None, None, 2, cs.ctypes.data_as(C.c_void_p), None, None, None, counts)
    assert rc == 0 and counts[0] == 0 and counts[1] == 0
    # and the library still works after the refusals
    assert int(one.sum().item()) == 0
