"""Row-partitioned SpMV over the GPUs of one box: one process per GPU (torch.distributed for the
plumbing), the operand x partitioned like the rows, remote ("halo") entries gathered by the product
kernel itself from the peers' HBM over NVLink (CUDA-IPC mappings; see DistX in csrc/slc_spmv.cu).

The reference has no distributed code; this is the multi-GPU form of mulV
(Data/Matrix/Sparse.hs:464-471) that BASELINE.json's north_star asks for.

Host-side pieces (partitioning, row blocks, handle exchange) are plain numpy / torch.distributed and
are covered by world_size-2 gloo tests on the CPU (tests/test_sharded_gloo.py).
"""
import ctypes as C
import os
import sys
import time

import numpy as np

from . import _capi
from .generators import CSC
from .matrix import Matrix, SparseLinearError, _check


IPC_BYTES = 72   # SLC_IPC_HANDLE_BYTES


# ---- partitioning (host) ---------------------------------------------------------------------------

def uniform_splits(n, nparts, align=1):
    """Contiguous blocks of (almost) equal size, boundaries on multiples of `align`."""
    units = -(-n // align)
    b = [min(n, align * ((units * p) // nparts)) for p in range(nparts)] + [n]
    return np.asarray(b, dtype=np.int64)


def balanced_splits(row_nnz, nparts):
    """Contiguous row blocks with (almost) equal numbers of nonzeros (power-law matrices)."""
    cum = np.concatenate([[0], np.cumsum(np.asarray(row_nnz, dtype=np.int64))])
    total = cum[-1]
    targets = (total * np.arange(1, nparts, dtype=np.int64)) // nparts
    inner = np.searchsorted(cum, targets, side="left")
    b = np.concatenate([[0], inner, [len(cum) - 1]]).astype(np.int64)
    return np.maximum.accumulate(b)


def row_block(m, r0, r1):
    """A[r0:r1, :] of a CSC matrix, as a CSC of shape (r1-r0) x ncols with local row indices."""
    keep = (m.rowidx >= r0) & (m.rowidx < r1)
    cols = np.repeat(np.arange(m.ncols, dtype=np.int64), np.diff(m.colptr))
    counts = np.bincount(cols[keep], minlength=m.ncols)
    colptr = np.concatenate([[0], np.cumsum(counts)]).astype(np.int32)
    return CSC(int(r1 - r0), m.ncols, colptr, (m.rowidx[keep] - r0).astype(np.int32),
               np.ascontiguousarray(m.values[keep]))


def exchange_objects(obj, dist, group=None):
    """all_gather of one small picklable object per rank (IPC handles, sizes)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return [obj]
    out = [None] * dist.get_world_size(group)
    dist.all_gather_object(out, obj, group=group)
    return out


def vstack_blocks(blocks):
    """Row blocks of one matrix (CSC, same ncols, listed top to bottom) put back together: column j of
    the result is the blocks' columns j one after the other, row indices shifted by the rows above.
    This is how the row-partitioned SpGEMM's outputs concatenate."""
    ncols = blocks[0].ncols
    counts = [np.diff(np.asarray(b.colptr, dtype=np.int64)) for b in blocks]
    colptr = np.concatenate([[0], np.cumsum(np.sum(counts, axis=0))])
    nnz = int(colptr[-1])
    rowidx = np.empty(nnz, dtype=np.int32)
    values = np.empty(nnz, dtype=np.asarray(blocks[0].values).dtype)
    before = np.zeros(ncols, dtype=np.int64)     # entries of column j already placed by the blocks above
    row0 = 0
    for b, cnt in zip(blocks, counts):
        n = int(cnt.sum())
        if n:
            src_col = np.repeat(np.arange(ncols, dtype=np.int64), cnt)
            within = np.arange(n, dtype=np.int64) - np.repeat(np.asarray(b.colptr[:-1], dtype=np.int64), cnt)
            dst = colptr[:-1][src_col] + before[src_col] + within
            rowidx[dst] = np.asarray(b.rowidx, dtype=np.int64) + row0
            values[dst] = b.values
        before += cnt
        row0 += b.nrows
    return CSC(row0, ncols, colptr.astype(np.int32), rowidx, values)


def mm_row_block(a, b, rank, world, splits=None):
    """This rank's share of C = A B (mm, Data/Matrix/Sparse.hs:691-702): rows [r0, r1) of A against a
    replicated B, no exchange while computing.  a: host CSC of the whole A (every rank can build or
    load it) -- only the block is uploaded; b: Matrix.  Returns (C block as Matrix, r0, r1)."""
    from .matrix import mm
    if splits is None:
        splits = uniform_splits(a.nrows, world)
    r0, r1 = int(splits[rank]), int(splits[rank + 1])
    blk = row_block(a, r0, r1)
    local = Matrix(blk.ncols, blk.nrows, blk.colptr, blk.rowidx, blk.values)
    return mm(local, b), r0, r1


def hstack_blocks(blocks):
    """Column blocks of one matrix (CSC, same nrows, listed left to right) side by side: the arrays are
    concatenated and the pointers shifted -- in CSC a column block is a contiguous slice, so this is the
    cheap direction (compare vstack_blocks)."""
    nrows = blocks[0].nrows
    ptrs, base = [np.zeros(1, dtype=np.int64)], 0
    for b in blocks:
        cp = np.asarray(b.colptr, dtype=np.int64)
        ptrs.append(cp[1:] + base)
        base += int(cp[-1])
    colptr = np.concatenate(ptrs)
    rowidx = np.concatenate([np.asarray(b.rowidx, dtype=np.int32) for b in blocks]) if blocks else np.zeros(0, np.int32)
    values = np.concatenate([np.asarray(b.values) for b in blocks])
    return CSC(nrows, int(len(colptr) - 1), colptr.astype(np.int32), rowidx, values)


def mm_col_block(a, b, rank, world, splits=None):
    """This rank's share of C = A B by COLUMNS of B: C[:, c0:c1] = A B[:, c0:c1] (mm is column by column,
    Data/Matrix/Sparse.hs:691-702, so a column block of C needs all of A and only that block of B).
    a: Matrix, the whole A resident on this rank; b: host CSC of the whole B -- only the block is
    uploaded.  Nothing is exchanged while computing and the blocks concatenate (hstack_blocks).
    Unlike mm_row_block every per-column pass of the product (bounds, classes, scans, pointers) shrinks
    with the block, which is what lets it scale: cfg 3 at 8 GPUs, see DESIGN.md section 5.
    Returns (C block as Matrix, c0, c1)."""
    from .matrix import mm
    if splits is None:
        splits = uniform_splits(b.ncols, world)
    c0, c1 = int(splits[rank]), int(splits[rank + 1])
    blk = col_block(b, c0, c1)
    local = Matrix(blk.ncols, blk.nrows, blk.colptr, blk.rowidx, blk.values)
    return mm(a, local), c0, c1


def gather_blocks(c_local, dist, group=None):
    """One gather of the C blocks when a single result is wanted (host side, rank order)."""
    blk = CSC(c_local.nrows, c_local.ncols, c_local.pointers, c_local.indices, c_local.values)
    return vstack_blocks(exchange_objects(blk, dist, group))


# ---- transpose and assembly over ranks: one exchange step, local kernels ----------
# Both are written as "make the pieces for every destination" / "combine what arrived", so a test can
# run the ranks one after the other in one process; exchange_pieces is the only collective.

def col_block(m, c0, c1):
    """A[:, c0:c1] of a CSC matrix (a slice of the arrays, pointers rebased)."""
    p0, p1 = int(m.colptr[c0]), int(m.colptr[c1])
    return CSC(m.nrows, int(c1 - c0), (np.asarray(m.colptr[c0:c1 + 1], dtype=np.int64) - p0).astype(np.int32),
               np.ascontiguousarray(m.rowidx[p0:p1]), np.ascontiguousarray(m.values[p0:p1]))


def exchange_pieces(send, dist, group=None):
    """All-to-all-v of picklable pieces: send[p] goes to rank p; returns recv with recv[q] = the piece
    rank q sent here.  Numpy arrays inside the pieces travel as NCCL all_to_all_single payloads when
    the default backend is NCCL (sizes first, then one flat byte buffer per pair); on gloo (the CPU
    tests) the pieces are gathered as objects."""
    if dist is None or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return [send[0]]
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    if dist.get_backend(group) != "nccl":
        everything = exchange_objects(send, dist, group)
        return [everything[q][rank] for q in range(world)]
    import pickle
    import torch
    blobs = [np.frombuffer(pickle.dumps(x, protocol=pickle.HIGHEST_PROTOCOL), dtype=np.uint8) for x in send]
    out_sizes = torch.tensor([len(b) for b in blobs], dtype=torch.int64, device="cuda")
    in_sizes = torch.empty(world, dtype=torch.int64, device="cuda")
    dist.all_to_all_single(in_sizes, out_sizes, group=group)
    in_sizes = [int(v) for v in in_sizes.tolist()]
    flat = torch.from_numpy(np.concatenate(blobs)).cuda()
    got = torch.empty(sum(in_sizes), dtype=torch.uint8, device="cuda")
    dist.all_to_all_single(got, flat, output_split_sizes=in_sizes, input_split_sizes=[len(b) for b in blobs],
                           group=group)
    got = got.cpu().numpy()
    recv, at = [], 0
    for n in in_sizes:
        recv.append(pickle.loads(got[at:at + n].tobytes()))
        at += n
    return recv


def transpose_pieces(local_t, row_splits):
    """local_t: the transpose of this rank's column block A[:, c0:c1], i.e. a (c1-c0) x nrows CSC whose
    column r is row r of the block.  Piece p = its columns [row_splits[p], row_splits[p+1])."""
    return [col_block(local_t, int(row_splits[p]), int(row_splits[p + 1])) for p in range(len(row_splits) - 1)]


def transpose_combine(received):
    """received[q] = piece from rank q (its rows are A's columns [c_q0, c_q1)); stacked in rank order they
    are this rank's column block of the transpose, A^T[:, r0:r1] -- stable, because rank order is column
    order of A (Sparse.hs:301-329 keeps equal keys in input order)."""
    return vstack_blocks(received)


def sharded_transpose(local_block, row_splits, dist=None, group=None, local_transpose=None):
    """Transpose of a matrix held as column blocks (rank q: A[:, c_q0:c_q1], all nrows rows).  Returns
    this rank's column block of A^T as a CSC of shape ncols x (r1-r0).  The local transpose runs on
    the GPU (slc_*_transpose); one all-to-all-v moves every entry to the rank that owns its row."""
    if local_transpose is None:
        def local_transpose(blk):
            from .matrix import transpose
            t = transpose(Matrix(blk.ncols, blk.nrows, blk.colptr, blk.rowidx, blk.values))
            return CSC(t.nrows, t.ncols, t.pointers, t.indices, t.values)
    pieces = transpose_pieces(local_transpose(local_block), row_splits)
    return transpose_combine(exchange_pieces(pieces, dist, group))


def compress_pieces(rows, cols, vals, col_splits):
    """Triplets bucketed by the rank that owns their column, input order kept inside a bucket, columns
    rebased to the owner's block.  Out-of-range columns go to the nearest block so that the owner's
    compress reports them (Sparse.hs:192-212)."""
    rows, cols, vals = np.asarray(rows), np.asarray(cols), np.asarray(vals)
    nparts = len(col_splits) - 1
    owner = np.clip(np.searchsorted(np.asarray(col_splits[1:-1]), cols, side="right"), 0, nparts - 1)
    out = []
    for p in range(nparts):
        keep = owner == p
        out.append((np.ascontiguousarray(rows[keep]), np.ascontiguousarray(cols[keep] - int(col_splits[p])),
                    np.ascontiguousarray(vals[keep])))
    return out


def compress_combine(nrows, width, received, local_compress=None):
    """Triplets from every rank, concatenated in rank order, assembled locally (slc_*_compress on the
    GPU): duplicates are summed in (source rank, input position) order."""
    if local_compress is None:
        from .matrix import compress as local_compress
    r = np.concatenate([x[0] for x in received])
    c = np.concatenate([x[1] for x in received])
    v = np.concatenate([x[2] for x in received])
    return local_compress(nrows, width, r, c, v)


def sharded_compress(nrows, rows, cols, vals, col_splits, rank, dist=None, group=None, local_compress=None):
    """compress (Sparse.hs:184-255) of triplets spread over the ranks: returns this rank's column block
    A[:, c0:c1] (nrows x (c1-c0)).  One all-to-all-v, then the local device assembly."""
    pieces = compress_pieces(rows, cols, vals, col_splits)
    width = int(col_splits[rank + 1] - col_splits[rank])
    return compress_combine(nrows, width, exchange_pieces(pieces, dist, group), local_compress)


# ---- the same two operations with every entry staying on the GPUs -----------------------------------------
# Local kernels (slc_*_transpose, slc_*_coo_bucket_dev, slc_*_compress_dev, slc_*_vstack_dev), pieces cut by
# slicing the handle's device arrays, ONE all-to-all-v of raw device buffers (NCCL all_to_all_single with
# split sizes; sizes first).  torch is plumbing here: views of device memory and the collective.

class _DevArray:
    """A raw device pointer as something torch.as_tensor can wrap without copying."""

    def __init__(self, ptr, n, typestr):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": typestr, "data": (int(ptr), False), "version": 2}


def _view(ptr, n, typestr):
    import torch
    if n == 0 or not ptr:
        return torch.empty(0, dtype={"<i4": torch.int32, "<f8": torch.float64}[typestr], device="cuda")
    return torch.as_tensor(_DevArray(ptr, n, typestr), device="cuda")


def device_arrays(m):
    """(colptr, rowidx, vals) of a Matrix's device handle as torch tensors viewing the handle's memory
    (Complex Double values come back as 2 float64 per entry).  The handle must outlive the views."""
    cp, ri, va = C.c_void_p(), C.c_void_p(), C.c_void_p()
    dm = m.device()
    _check(_capi.lib().slc_device_arrays(dm.handle, C.byref(cp), C.byref(ri), C.byref(va)))
    nr, nc, nnz = dm.dims()
    w = 2 if m._p == "z" else 1
    return _view(cp.value, nc + 1, "<i4"), _view(ri.value, nnz, "<i4"), _view(va.value, nnz * w, "<f8")


class _Trace:
    """SLC_SHARD_TRACE=1: wall time of every phase of the sharded transpose / assembly (each closed by a device
    synchronisation) on stderr, rank 0 only."""

    def __init__(self, what, dist):
        self.on = os.environ.get("SLC_SHARD_TRACE", "0") != "0" and (dist is None or not dist.is_initialized() or dist.get_rank() == 0)
        self.what, self.t = what, time.perf_counter()

    def mark(self, phase):
        if not self.on:
            return
        import torch
        torch.cuda.synchronize()
        now = time.perf_counter()
        sys.stderr.write("[slc %s] %-28s %8.3f ms\n" % (self.what, phase, (now - self.t) * 1e3))
        self.t = now


def _all_to_all_v(send, send_counts, dist, group, width=1, recv_counts=None):
    """send: 1-D device tensor laid out bucket after bucket; send_counts[p] elements (x width) go to rank p.
    recv_counts: what every rank sends here, when the caller already knows it (several arrays cut the same
    way need the sizes exchanged once, not once per array).  Returns (received tensor, recv_counts)."""
    import torch
    world = len(send_counts)
    if dist is None or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return send, list(send_counts)
    if recv_counts is None:
        sc = torch.tensor(send_counts, dtype=torch.int64, device="cuda")
        rc = torch.empty(world, dtype=torch.int64, device="cuda")
        dist.all_to_all_single(rc, sc, group=group)
        recv_counts = [int(v) for v in rc.tolist()]
    out = torch.empty(sum(recv_counts) * width, dtype=send.dtype, device="cuda")
    dist.all_to_all_single(out, send, output_split_sizes=[c * width for c in recv_counts],
                           input_split_sizes=[c * width for c in send_counts], group=group)
    return out, recv_counts


def vstack_device(p, block_rows, ncols, ptrs, idxs, vals):
    """slc_*_vstack_dev over lists of device tensors (one CSC block each, top to bottom)."""
    n = len(block_rows)
    br = np.ascontiguousarray(block_rows, dtype=np.int32)
    pp = (C.c_void_p * n)(*[t.data_ptr() for t in ptrs])
    ii = (C.c_void_p * n)(*[t.data_ptr() if t.numel() else None for t in idxs])
    vv = (C.c_void_p * n)(*[t.data_ptr() if t.numel() else None for t in vals])
    h = C.c_void_p()
    _check(getattr(_capi.lib(), "slc_%s_vstack_dev" % p)(n, br.ctypes.data_as(C.c_void_p), int(ncols), C.cast(pp, C.c_void_p),
                                                        C.cast(ii, C.c_void_p), C.cast(vv, C.c_void_p), C.byref(h)))
    return Matrix._from_device(h)


def sharded_transpose_device(local, row_splits, dist=None, group=None, col_splits=None):
    """Transpose of a matrix held as column blocks (this rank: `local` = A[:, c0:c1] as a Matrix), entries
    never leaving the GPUs.  Returns this rank's column block of A^T, ncols(A) x (r1-r0), as a Matrix.
    Stable (Sparse.hs:301-329): the pieces are stacked in rank order = A's column order."""
    import torch
    from .matrix import transpose
    world = len(row_splits) - 1
    rank = dist.get_rank(group) if (dist is not None and dist.is_initialized() and world > 1) else 0
    if col_splits is None:
        col_splits = row_splits
    w = 2 if local._p == "z" else 1
    tr = _Trace("sharded transpose", dist)
    lt = transpose(local)                                   # (c1-c0) x nrows: column r = row r of the block
    tr.mark("local transpose")
    cp, ri, va = device_arrays(lt)
    cuts = torch.as_tensor(np.asarray(row_splits, dtype=np.int64), device="cuda")
    bounds = [int(v) for v in cp[cuts].tolist()]            # world+1 pointer values
    tr.mark("cut points")
    send_counts = [bounds[q + 1] - bounds[q] for q in range(world)]
    colcnt = cp[1:] - cp[:-1]                               # entries per column of local^T, laid out piece after piece
    widths = [int(row_splits[q + 1] - row_splits[q]) for q in range(world)]
    np_cols = widths[rank]
    if world == 1:
        rcnt, ridx, rval, recv_counts = colcnt, ri, va, send_counts
    else:
        rcnt = torch.empty(np_cols * world, dtype=torch.int32, device="cuda")
        dist.all_to_all_single(rcnt, colcnt.contiguous(), output_split_sizes=[np_cols] * world, input_split_sizes=widths, group=group)
        tr.mark("exchange of column counts")
        # what every rank sends here follows from the counts just received: no second size exchange
        recv_counts = [int(v) for v in rcnt.view(world, np_cols).sum(dim=1, dtype=torch.int64).tolist()]
        ridx, _ = _all_to_all_v(ri, send_counts, dist, group, recv_counts=recv_counts)
        tr.mark("exchange of indices")
        rval, _ = _all_to_all_v(va, send_counts, dist, group, width=w, recv_counts=recv_counts)
        tr.mark("exchange of values")
    # per-piece pointers: exclusive scan of the per-column counts of each piece
    # (one flat scan and a subtraction: a scan along the long side of a (world x np_cols) tensor runs one block per
    # row in torch -- 1.7 ms for 2 x 1 M counts)
    flat = torch.cumsum(rcnt, dim=0, dtype=torch.int32).view(world, np_cols)
    before = torch.zeros(world, dtype=torch.int32, device="cuda")
    before[1:] = flat[:-1, -1]
    ptr = torch.zeros((world, np_cols + 1), dtype=torch.int32, device="cuda")
    ptr[:, 1:] = flat - before[:, None]
    offs = np.concatenate([[0], np.cumsum(recv_counts)]).astype(np.int64)
    tr.mark("piece pointers")
    out = vstack_device(local._p, [int(col_splits[q + 1] - col_splits[q]) for q in range(world)], np_cols,
                        [ptr[q] for q in range(world)], [ridx[offs[q]:offs[q + 1]] for q in range(world)],
                        [rval[offs[q] * w:offs[q + 1] * w] for q in range(world)])
    tr.mark("stacking")
    del lt
    return out


def sharded_compress_device(nrows, rows_dev, cols_dev, vals_dev, col_splits, rank, dist=None, group=None, complex_values=False):
    """compress (Sparse.hs:184-255) of triplets spread over the ranks and resident in HBM (torch int32 /
    int32 / float64 tensors; complex values as 2 float64 per entry): returns this rank's column block
    A[:, c0:c1] as a Matrix.  Bucket by owner on the device (input order kept), one all-to-all-v of the
    three arrays, local device assembly (duplicates summed in (source rank, input position) order)."""
    import torch
    from .matrix import compress_dev
    world = len(col_splits) - 1
    p = "z" if complex_values else "d"
    w = 2 if complex_values else 1
    n = int(rows_dev.numel())
    cs = np.ascontiguousarray(col_splits, dtype=np.int32)
    tr = _Trace("sharded assembly", dist)
    orow, ocol, oval = torch.empty_like(rows_dev), torch.empty_like(cols_dev), torch.empty_like(vals_dev)
    counts = (C.c_int64 * world)()
    _check(getattr(_capi.lib(), "slc_%s_coo_bucket_dev" % p)(
        n, C.c_void_p(rows_dev.data_ptr()), C.c_void_p(cols_dev.data_ptr()), C.c_void_p(vals_dev.data_ptr()), world,
        cs.ctypes.data_as(C.c_void_p), C.c_void_p(orow.data_ptr()), C.c_void_p(ocol.data_ptr()), C.c_void_p(oval.data_ptr()), counts))
    send_counts = [int(c) for c in counts]
    tr.mark("bucket by owner")
    rrow, recv_counts = _all_to_all_v(orow, send_counts, dist, group)
    tr.mark("sizes + exchange of rows")
    rcol, _ = _all_to_all_v(ocol, send_counts, dist, group, recv_counts=recv_counts)
    rval, _ = _all_to_all_v(oval, send_counts, dist, group, width=w, recv_counts=recv_counts)
    tr.mark("exchange of columns, values")
    width = int(cs[rank + 1] - cs[rank])
    total = sum(recv_counts)
    out = compress_dev(nrows, width, total, rrow.data_ptr() if total else 0, rcol.data_ptr() if total else 0,
                       rval.data_ptr() if total else 0, np.complex128 if complex_values else np.float64)
    tr.mark("local assembly")
    return out


# ---- device side --------------------------------------------------------------------------------------

class ShardedSpMV:
    """y_local = A[r0:r1, :] x with x sliced by `splits` over the ranks (one process per GPU).

    local: Matrix of shape (r1-r0) x ncols held by this rank.  Each rank owns x[splits[r]:splits[r+1]]
    in TWO cudaMalloc'ed buffers that every other rank maps through CUDA IPC; products alternate
    between them (set_x writes the buffer the NEXT product reads), which is what makes rewriting x
    every step safe without a second handshake (include/slc.h, "Ordering contract").
    The ranks' dependency sets (who reads whose slice) are exchanged once, so a product waits for its
    neighbours only, not for every rank.
    """

    def __init__(self, local, splits, rank, world, dist=None, cpu_group=None, k=None, mode="halo"):
        self.local, self.rank, self.world, self.dist = local, rank, world, dist
        self.k = k                      # None: vectors (mulV); k >= 1: column-major blocks (mulM)
        # "halo": remote entries are gathered by the product kernel itself (banded matrices);
        # "allgather": every slice is pulled over NVLink into a full local copy first (power-law)
        if mode not in ("halo", "allgather") or (mode == "allgather" and k is not None):
            raise SparseLinearError("ShardedSpMV: unsupported mode %r" % (mode,))
        self.mode = mode
        ncol = 1 if k is None else int(k)
        self.splits = np.ascontiguousarray(splits, dtype=np.int32)
        if len(self.splits) != world + 1 or self.splits[0] != 0 or self.splits[-1] != local.ncols:
            raise SparseLinearError("axpy_: column dimension %d does not match operand dimension %d"
                                    % (local.ncols, int(self.splits[-1])))
        self.p = local._p
        self.el = 16 if self.p == "z" else 8
        lib = _capi.lib()
        self.nloc_x = int(self.splits[rank + 1] - self.splits[rank])
        # two operand buffers carved from ONE allocation (one IPC handle per rank; see slc_ipc_export)
        self.x_bytes = -(-max(self.nloc_x * ncol, 1) * self.el // 256) * 256
        self.x_base = C.c_void_p()
        _check(lib.slc_dev_alloc(C.byref(self.x_base), 2 * self.x_bytes))
        self.x_bufs = [self.x_base.value, self.x_base.value + self.x_bytes]
        self.cur = 0                    # buffer the next product reads is x_bufs[cur ^ 1] until set_x / product flips
        self.y_ptr = C.c_void_p()
        _check(lib.slc_dev_alloc(C.byref(self.y_ptr), max(local.nrows * ncol, 1) * self.el))
        self.xfull_ptr = C.c_void_p()
        if mode == "allgather":
            _check(lib.slc_dev_alloc(C.byref(self.xfull_ptr), max(local.ncols, 1) * self.el))
        handle = C.create_string_buffer(IPC_BYTES)
        _check(lib.slc_ipc_export(self.x_base, handle))
        self.group = C.c_void_p()
        ghandle = C.create_string_buffer(IPC_BYTES)
        _check(lib.slc_peer_group_create(world, rank, C.byref(self.group), ghandle))
        self.local.device().prepare()
        need = C.c_uint32(0)
        _check(getattr(lib, "slc_%s_dist_needs" % self.p)(self.local.device().handle, world, rank,
                                                          self.splits.ctypes.data_as(C.c_void_p), C.byref(need)))
        if mode == "allgather":
            need = C.c_uint32((1 << world) - 1)
        both = exchange_objects((bytes(handle.raw), bytes(ghandle.raw), int(need.value)), dist, cpu_group)
        handles = [b[0] for b in both]
        _check(lib.slc_peer_group_connect(self.group, C.create_string_buffer(b"".join(b[1] for b in both),
                                                                            IPC_BYTES * world)))
        self.deps = dependency_mask([b[2] for b in both], rank)
        _check(lib.slc_peer_group_set_deps(self.group, self.deps))
        self.use_nccl_barrier = os.environ.get("SLC_BARRIER", "peer") == "nccl"
        self.tables = [(C.c_void_p * world)(), (C.c_void_p * world)()]    # part pointers of buffer 0 / 1
        self._imported = []
        for r in range(world):
            if r == rank:
                base = self.x_base.value
                xb = self.x_bytes
            else:
                ptr = C.c_void_p()
                _check(lib.slc_ipc_import(C.create_string_buffer(handles[r], IPC_BYTES), C.byref(ptr)))
                self._imported.append(ptr)
                base = ptr.value
                nloc = int(self.splits[r + 1] - self.splits[r])
                xb = -(-max(nloc * ncol, 1) * self.el // 256) * 256
            self.tables[0][r] = base
            self.tables[1][r] = base + xb
        self._fn = getattr(lib, "slc_%s_%s_dist_dev" % (self.p, "mulv" if k is None else "mulm"))
        self._fn_ag = getattr(lib, "slc_%s_mulv_allgather_dev" % self.p)
        # every part's block is packed: leading dimension = its own row count
        self.lds = np.ascontiguousarray(np.diff(self.splits).astype(np.int64))
        self._flag = None

    @property
    def x_ptr(self):
        """Device address of the buffer the next product reads (fill it, then call product / mulv)."""
        return C.c_void_p(self.x_bufs[self.cur ^ 1])

    def set_x(self, x_local):
        """This rank's slice of x (vector) or its row block of X (nloc x k, stored column-major) for the
        NEXT product: written into the buffer the previous product did not read."""
        if self.k is None:
            x_local = np.ascontiguousarray(x_local, dtype=self.local.dtype)
            assert len(x_local) == self.nloc_x
        else:
            x_local = np.asfortranarray(x_local, dtype=self.local.dtype)
            assert x_local.shape == (self.nloc_x, self.k)
        _check(_capi.lib().slc_memcpy_h2d(self.x_ptr, C.c_void_p(x_local.ctypes.data), x_local.nbytes))

    def barrier(self):
        """Order this rank's next product after every peer's writes to its x slice.  Enqueued on the
        current stream, no host synchronisation: the library's own peer-memory barrier kernel
        (csrc/slc_peer.cu), or with SLC_BARRIER=nccl a one-element NCCL all-reduce."""
        if self.world == 1:
            return
        if self.use_nccl_barrier:
            import torch
            if self._flag is None:
                self._flag = torch.zeros(1, device="cuda")
            self.dist.all_reduce(self._flag)
        else:
            _check(_capi.lib().slc_peer_barrier(self.group))

    def barrier_all(self):
        """Full rendezvous of all ranks' streams (every rank waits for every rank), on the device."""
        if self.world > 1:
            _check(_capi.lib().slc_peer_barrier_all(self.group))

    def timed_out(self):
        f = C.c_int(0)
        _check(_capi.lib().slc_peer_group_timed_out(self.group, C.byref(f)))
        return bool(f.value)

    def stats(self, reset=False):
        """(epochs completed, timed waits, ns waited in total, longest wait in ns); synchronises."""
        out = (C.c_int64 * 4)()
        _check(_capi.lib().slc_peer_group_stats(self.group, out, 1 if reset else 0))
        return tuple(int(v) for v in out)

    def product(self, sync=None):
        """One product over the buffer set_x / x_ptr last designated; flips the buffers."""
        self.cur ^= 1
        table = self.tables[self.cur]
        if self.mode == "allgather":
            _check(self._fn_ag(self.local.device().handle, self.world, self.rank,
                               self.splits.ctypes.data_as(C.c_void_p), C.cast(table, C.c_void_p),
                               self.xfull_ptr, self.y_ptr, sync))
        elif self.k is None:
            _check(self._fn(self.local.device().handle, self.world, self.rank,
                            self.splits.ctypes.data_as(C.c_void_p), C.cast(table, C.c_void_p), self.y_ptr,
                            sync))
        else:
            _check(self._fn(self.local.device().handle, self.world, self.rank,
                            self.splits.ctypes.data_as(C.c_void_p), C.cast(table, C.c_void_p),
                            self.lds.ctypes.data_as(C.c_void_p), self.y_ptr, max(self.local.nrows, 1), self.k,
                            sync))

    def repeat_x(self):
        """Keep the operand for another product without re-sending it: copy the current buffer into the
        other one on the device (benchmarks that time products of a fixed x)."""
        lib = _capi.lib()
        src, dst = self.x_bufs[self.cur], self.x_bufs[self.cur ^ 1]
        _check(lib.slc_memcpy_d2d(C.c_void_p(dst), C.c_void_p(src), self.x_bytes))

    def mulv(self):
        """One distributed product, ordered against the peers.  Default: the library signals, runs
        the rows that need no peer, waits, then runs the boundary rows (the peers' latency hides
        behind the interior).  SLC_BARRIER=nccl: a full NCCL barrier first, then the product."""
        if self.world == 1 or self.use_nccl_barrier:
            self.barrier()
            self.product(None)
        else:
            self.product(self.group)

    mulm = mulv   # the block form: same ordering, slc_*_mulm_dist_dev underneath

    def y(self):
        shape = self.local.nrows if self.k is None else (self.local.nrows, self.k)
        out = np.empty(shape, dtype=self.local.dtype, order="F")
        _check(_capi.lib().slc_memcpy_d2h(C.c_void_p(out.ctypes.data), self.y_ptr, out.nbytes))
        return out

    def close(self):
        lib = _capi.lib()
        for p in self._imported:
            lib.slc_ipc_close(p)
        self._imported = []
        if self.group:
            lib.slc_peer_group_free(self.group)
            self.group = None
        if self.x_base:
            lib.slc_dev_free(self.x_base)
            self.x_base = None
        if self.y_ptr:
            lib.slc_dev_free(self.y_ptr)
            self.y_ptr = None
        if self.xfull_ptr:
            lib.slc_dev_free(self.xfull_ptr)
            self.xfull_ptr = None


def dependency_mask(needs, rank):
    """needs[q] = bit mask of the parts rank q reads from.  Rank `rank` depends on the parts it reads
    and on the ranks that read it (they must be done with a buffer before it is rewritten)."""
    deps = int(needs[rank])
    for q, nq in enumerate(needs):
        if (int(nq) >> rank) & 1:
            deps |= 1 << q
    return deps & ~(1 << rank)


class ShardedMatrix:
    """One process, several GPUs: slc_{d,z}_upload_sharded / mulv_sharded / mulm_sharded (include/slc.h).
    The host-side mirror of what the Haskell shim binds for the multi-GPU products."""

    def __init__(self, m, ndev, devices=None):
        lib = _capi.lib()
        self.p = m._p
        self.dtype = m.dtype
        self.nrows, self.ncols = m.nrows, m.ncols
        self.h = C.c_void_p()
        dv = None if devices is None else np.ascontiguousarray(devices, dtype=np.int32)
        _check(getattr(lib, "slc_%s_upload_sharded" % self.p)(
            int(ndev), None if dv is None else dv.ctypes.data_as(C.c_void_p), m.nrows, m.ncols,
            C.c_void_p(m.pointers.ctypes.data), C.c_void_p(m.indices.ctypes.data), C.c_void_p(m.values.ctypes.data),
            C.byref(self.h)))
        self.ndev = int(ndev)

    def splits(self):
        rs = np.zeros(self.ndev + 1, dtype=np.int32)
        cs = np.zeros(self.ndev + 1, dtype=np.int32)
        _check(_capi.lib().slc_sharded_splits(self.h, rs.ctypes.data_as(C.c_void_p), cs.ctypes.data_as(C.c_void_p)))
        return rs, cs

    def mulV(self, x):
        x = np.ascontiguousarray(x, dtype=self.dtype)
        if x.ndim != 1 or len(x) != self.ncols:
            raise SparseLinearError("axpy_: column dimension %d does not match operand dimension %d" % (self.ncols, x.size))
        y = np.empty(self.nrows, dtype=self.dtype)
        _check(getattr(_capi.lib(), "slc_%s_mulv_sharded" % self.p)(self.h, C.c_void_p(x.ctypes.data), C.c_void_p(y.ctypes.data)))
        return y

    def mulM(self, X):
        X = np.asarray(X)
        if X.ndim != 2 or X.shape[0] != self.ncols:
            raise SparseLinearError("mulM: inner dimension mismatch")
        Xf = np.asfortranarray(X, dtype=self.dtype)
        k = Xf.shape[1]
        Y = np.empty((self.nrows, k), dtype=self.dtype, order="F")
        _check(getattr(_capi.lib(), "slc_%s_mulm_sharded" % self.p)(
            self.h, C.c_void_p(Xf.ctypes.data), max(self.ncols, 1), C.c_void_p(Y.ctypes.data), max(self.nrows, 1), k))
        return Y

    def close(self):
        if self.h:
            _capi.lib().slc_sharded_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
