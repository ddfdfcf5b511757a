"""Row-partitioned SpMV over the GPUs of one box: one process per GPU (torch.distributed for the
plumbing), the operand x partitioned like the rows, remote ("halo") entries gathered by the product
kernel itself from the peers' HBM over NVLink (CUDA-IPC mappings; see DistX in csrc/slc_spmv.cu).

The reference has no distributed code; this is the multi-GPU form of mulV
(Data/Matrix/Sparse.hs:464-471) that BASELINE.json's north_star asks for.

Host-side pieces (partitioning, row blocks, handle exchange) are plain numpy / torch.distributed and
are covered by world_size-2 gloo tests on the CPU (tests/test_sharded_gloo.py).
"""
import ctypes as C
import json
import os
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


# ---- device side --------------------------------------------------------------------------------------

class ShardedSpMV:
    """y_local = A[r0:r1, :] x with x sliced by `splits` over the ranks.

    local: Matrix of shape (r1-r0) x ncols held by this rank.  Each rank owns x[splits[r]:splits[r+1]]
    in a cudaMalloc'ed buffer that every other rank maps through CUDA IPC.
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
        self.x_ptr = C.c_void_p()
        _check(lib.slc_dev_alloc(C.byref(self.x_ptr), max(self.nloc_x * ncol, 1) * self.el))
        self.y_ptr = C.c_void_p()
        _check(lib.slc_dev_alloc(C.byref(self.y_ptr), max(local.nrows * ncol, 1) * self.el))
        self.xfull_ptr = C.c_void_p()
        if mode == "allgather":
            _check(lib.slc_dev_alloc(C.byref(self.xfull_ptr), max(local.ncols, 1) * self.el))
        handle = C.create_string_buffer(IPC_BYTES)
        _check(lib.slc_ipc_export(self.x_ptr, handle))
        self.group = C.c_void_p()
        ghandle = C.create_string_buffer(IPC_BYTES)
        _check(lib.slc_peer_group_create(world, rank, C.byref(self.group), ghandle))
        both = exchange_objects((bytes(handle.raw), bytes(ghandle.raw)), dist, cpu_group)
        handles = [b[0] for b in both]
        _check(lib.slc_peer_group_connect(self.group, C.create_string_buffer(b"".join(b[1] for b in both),
                                                                            IPC_BYTES * world)))
        self.use_nccl_barrier = os.environ.get("SLC_BARRIER", "peer") == "nccl"
        self.parts = (C.c_void_p * world)()
        self._imported = []
        for r in range(world):
            if r == rank:
                self.parts[r] = self.x_ptr.value
            else:
                ptr = C.c_void_p()
                _check(lib.slc_ipc_import(C.create_string_buffer(handles[r], IPC_BYTES), C.byref(ptr)))
                self.parts[r] = ptr.value
                self._imported.append(ptr)
        self.local.device().prepare()
        self._fn = getattr(lib, "slc_%s_%s_dist_dev" % (self.p, "mulv" if k is None else "mulm"))
        self._fn_ag = getattr(lib, "slc_%s_mulv_allgather_dev" % self.p)
        # every part's block is packed: leading dimension = its own row count
        self.lds = np.ascontiguousarray(np.diff(self.splits).astype(np.int64))
        self._flag = None

    def set_x(self, x_local):
        """This rank's slice of x (vector) or its row block of X (nloc x k, stored column-major)."""
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

    def timed_out(self):
        f = C.c_int(0)
        _check(_capi.lib().slc_peer_group_timed_out(self.group, C.byref(f)))
        return bool(f.value)

    def product(self, sync=None):
        if self.mode == "allgather":
            _check(self._fn_ag(self.local.device().handle, self.world, self.rank,
                               self.splits.ctypes.data_as(C.c_void_p), C.cast(self.parts, C.c_void_p),
                               self.xfull_ptr, self.y_ptr, sync))
        elif self.k is None:
            _check(self._fn(self.local.device().handle, self.world, self.rank,
                            self.splits.ctypes.data_as(C.c_void_p), C.cast(self.parts, C.c_void_p), self.y_ptr,
                            sync))
        else:
            _check(self._fn(self.local.device().handle, self.world, self.rank,
                            self.splits.ctypes.data_as(C.c_void_p), C.cast(self.parts, C.c_void_p),
                            self.lds.ctypes.data_as(C.c_void_p), self.y_ptr, max(self.local.nrows, 1), self.k,
                            sync))

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
        if self.x_ptr:
            lib.slc_dev_free(self.x_ptr)
            self.x_ptr = None
        if self.y_ptr:
            lib.slc_dev_free(self.y_ptr)
            self.y_ptr = None
        if self.xfull_ptr:
            lib.slc_dev_free(self.xfull_ptr)
            self.xfull_ptr = None


# ---- bench.py --gpus N ------------------------------------------------------------------------------------

def bench_main(args, rank, world, local_rank, dist):
    """cfg 5 row-partitioned over `world` GPUs.  strong: the 200^3 grid is cut into z-slabs;
    weak: every rank holds its own 200-plane slab of a (200*world) x 200 x 200 grid."""
    import torch
    import sparse_linear_b200 as sl
    from . import generators as g
    import bench as B

    if args.workload == "cfg3":
        return bench_spgemm_sharded(args, rank, world, local_rank, dist)
    if args.workload == "cfg4":
        return bench_powerlaw_sharded(args, rank, world, local_rank, dist)
    if args.workload in ("cfg3t", "cfg3a"):
        return bench_structural_sharded(args, rank, world, local_rank, dist)
    n = args.grid or 200
    if args.scaling == "weak":
        shape = (n * world, n, n)
    else:
        shape = (n, n, n)
    plane = shape[1] * shape[2]
    ntot = shape[0] * plane
    splits = uniform_splits(shape[0], world) * plane        # z-slabs
    r0, r1 = int(splits[rank]), int(splits[rank + 1])
    stream = torch.cuda.current_stream()
    sl.set_stream(stream.cuda_stream)
    cpu_group = dist.new_group(backend="gloo") if world > 1 else None

    # A is symmetric, so the row block A[r0:r1, :] is the transpose of the column block A[:, r0:r1]
    colblock = g.stencil3d_27pt(shape, col_range=(r0, r1))
    cb = Matrix(colblock.ncols, colblock.nrows, colblock.colptr, colblock.rowidx, colblock.values)
    local = sl.transpose(cb)                                 # (r1-r0) x ntot, on the device
    del cb
    k = 64 if args.workload == "cfg5mm" else None
    sh = ShardedSpMV(local, splits, rank, world, dist, cpu_group, k=k)
    xfull = g.uniform_vector(ntot, seed=0)
    nnz_local = int(colblock.colptr[-1])
    if k is None:
        sh.set_x(xfull[r0:r1])
        abytes_local = g.spmv_bytes(r1 - r0, r1 - r0, nnz_local)   # every x entry is read from HBM once, by its owner
    else:
        # column j of the block is x scaled by 1 + j/64, so one host SpMV checks all 64 columns
        colscale = 1.0 + np.arange(k) / 64.0
        sh.set_x(np.asfortranarray(xfull[r0:r1, None] * colscale[None, :]))
        abytes_local = g.spmm_bytes(r1 - r0, r1 - r0, nnz_local, k)
    tot = torch.tensor([float(abytes_local), float(nnz_local)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tot)
    abytes, nnz = float(tot[0].item()), int(tot[1].item())

    for _ in range(max(args.warmup, 3)):
        sh.mulv()
    torch.cuda.synchronize()
    # parity on this rank's rows: y_local = (column block)^T x, summed on the host
    want = np.add.reduceat(colblock.values * xfull[colblock.rowidx], colblock.colptr[:-1].astype(np.int64))
    got = sh.y()
    if k is not None:
        want = want[:, None] * colscale[None, :]
    err = float(np.linalg.norm(got - want) / max(np.linalg.norm(want), 1e-300))

    sampler = B.ClockSampler(local_rank)
    l0 = sl.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler.start()
    e0.record(stream)
    for _ in range(args.steps):
        sh.mulv()
    e1.record(stream)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    clocks = sampler.stop()
    launches = sl.launch_count() - l0
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    errt = torch.tensor([err], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(errt, op=dist.ReduceOp.MAX)
    ms = float(t.item()) / args.steps
    value = abytes / (ms * 1e-3) / 1e9

    # end to end: every step moves this rank's x slice from pinned host memory, multiplies, and
    # brings its y slice back
    e2e = None
    if not args.no_e2e:
        lib = _capi.lib()
        ncol = 1 if k is None else k
        if k is None:
            xp = torch.from_numpy(np.ascontiguousarray(xfull[r0:r1])).pin_memory()
        else:
            xp = torch.from_numpy(np.ascontiguousarray((xfull[r0:r1, None] * colscale[None, :]).T)).pin_memory()
        yp = torch.empty((r1 - r0) * ncol, dtype=torch.float64).pin_memory()
        steps_e2e = max(3, min(args.steps, 20 if k is None else 5))

        def hstep():
            _check(lib.slc_memcpy_h2d(sh.x_ptr, C.c_void_p(xp.data_ptr()), xp.numel() * 8))
            sh.mulv()
            _check(lib.slc_memcpy_d2h(C.c_void_p(yp.data_ptr()), sh.y_ptr, yp.numel() * 8))
        for _ in range(2):
            hstep()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(steps_e2e):
            hstep()
        torch.cuda.synchronize()
        dt = torch.tensor([(time.perf_counter() - t0) / steps_e2e], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        dt = float(dt.item())
        e2e = {"value": abytes / dt / 1e9, "unit": "GB/s", "ms_per_step": dt * 1e3, "steps": steps_e2e,
               "h2d_bytes_per_step": int(ntot * 8 * ncol), "d2h_bytes_per_step": int(ntot * 8 * ncol),
               "api": "ShardedSpMV (slc_memcpy_h2d + slc_d_%s_dist_dev + slc_memcpy_d2h per rank)"
                      % ("mulv" if k is None else "mulm")}

    if rank == 0:
        peak, peak_src = B.peaks()
        per_gpu = value / world
        line = {
            "metric": B.METRIC if k is None else "spmm_effective_hbm_gbps", "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic (seeded numpy PCG64)",
            "config": {"workload": ("cfg5: 3D 27-point stencil %dx%dx%d, Double SpMV, row-partitioned in z-slabs over %d GPU(s); "
                                    "halo entries gathered from peer HBM over NVLink inside the product kernel; "
                                    "the inter-GPU barrier (peer-memory flags) is fused into the same launch" % (shape + (world,)))
                       if k is None else
                       ("cfg5: 3D 27-point stencil %dx%dx%d, Double SpMM k=64, row-partitioned in z-slabs over %d GPU(s); "
                        "signal, interior rows (k_spmm_panel), wait, boundary rows gathering the peers' rows of X "
                        "over NVLink (k_spmm_boundary)" % (shape + (world,))),
                       "rows": ntot, "nnz": nnz, "algorithmic_bytes_per_step": abytes,
                       "plan": local.device().plan_info(),
                       "l2": "inputs larger than L2 (%.0f MB per GPU per step)" % (abytes / world / 1e6),
                       "pct_of_nominal_8TBs_per_gpu": 100.0 * per_gpu / B.NOMINAL_HBM_GBS,
                       "max_rel_err_vs_host_sum": float(errt.item())},
            "clocks": clocks, "gpu_launches": int(launches), "e2e": e2e,
            "roofline": {"bound": "hbm", "achieved": per_gpu, "peak": peak, "unit": "GB/s", "frac": per_gpu / peak,
                         "traffic": None, "peak_source": peak_src,
                         "kernel": "k_spmv_sell_dist" if k is None else "k_spmm_panel + k_spmm_boundary",
                         "note": "per-GPU figure; step time includes the inter-GPU barrier",
                         "barrier": "nccl all-reduce" if sh.use_nccl_barrier else
                         ("peer-memory flags, fused in k_spmv_sell_dist" if k is None else "peer-memory flags (signal / wait kernels)"),
                         "barrier_timed_out": sh.timed_out()},
            "cpu_baseline": None,
        }
        print(json.dumps(line), flush=True)
    sh.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def _max_over_ranks(ms, dist, world):
    import torch
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def _sum_over_ranks(vals, dist, world):
    import torch
    t = torch.tensor([float(v) for v in vals], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t)
    return [float(v) for v in t.tolist()]


def bench_spgemm_sharded(args, rank, world, local_rank, dist):
    """cfg 3, C = A A, strong-scaled: A cut into row blocks, B = A replicated, no exchange while
    computing; value = output nnz of all blocks / slowest rank's time (symbolic + numeric)."""
    import torch
    import sparse_linear_b200 as sl
    from . import generators as g
    import bench as B
    n = args.grid or 128
    a = g.laplace3d_7pt(n)
    stream = torch.cuda.current_stream()
    sl.set_stream(stream.cuda_stream)
    bm = Matrix(a.ncols, a.nrows, a.colptr, a.rowidx, a.values)
    splits = uniform_splits(n, world) * (n * n)           # z-slabs: equal work per rank
    r0, r1 = int(splits[rank]), int(splits[rank + 1])
    blk = row_block(a, r0, r1)
    local = Matrix(blk.ncols, blk.nrows, blk.colptr, blk.rowidx, blk.values)
    c = sl.mm(local, bm)
    sl.synchronize()
    nnzc_local = sl.nonZero(c)
    # parity of this rank's block against the closed form of the whole product and, on a small
    # grid, the structure of a host product (scipy is a second opinion only; tests use the oracle)
    del c
    steps = max(1, min(args.steps, 10))
    for _ in range(2):
        c = sl.mm(local, bm)
        del c
    sampler = B.ClockSampler(local_rank)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    l0 = sl.launch_count()
    sampler.start()
    e0.record(stream)
    for _ in range(steps):
        c = sl.mm(local, bm)
        del c
    e1.record(stream)
    torch.cuda.synchronize()
    clocks = sampler.stop()
    launches = sl.launch_count() - l0
    ms = _max_over_ranks(e0.elapsed_time(e1) / steps, dist, world)
    nnzc, nnza = _sum_over_ranks([nnzc_local, blk.colptr[-1]], dist, world)
    if rank == 0:
        peak, peak_src = B.peaks()
        nnzb = int(a.colptr[-1])
        floor_bytes = (nnza + nnzb * world + nnzc) * 12 + 3 * (a.ncols + 1) * 4 * world
        gbs = floor_bytes / (ms * 1e-3) / 1e9 / world
        expect = None
        if n >= 5:   # nnz(A A) of the 7-point Laplacian on an n^3 grid: the 25-point radius-2 diamond, clipped
            import itertools
            offs = {(dx + ex, dy + ey, dz + ez) for (dx, dy, dz) in [(0, 0, 0), (1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0), (0, 0, 1), (0, 0, -1)]
                    for (ex, ey, ez) in [(0, 0, 0), (1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0), (0, 0, 1), (0, 0, -1)]}
            expect = sum((n - abs(dx)) * (n - abs(dy)) * (n - abs(dz)) for dx, dy, dz in offs)
        line = {"metric": "spgemm_output_nnz_per_s", "value": nnzc / (ms * 1e-3), "unit": "nnz/s", "n_gpus": world,
                "steps": steps, "warmup": 2, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": "cfg3: SpGEMM A*A, 3D 7-point Laplacian %d^3, A in %d row blocks (z-slabs), "
                                       "B replicated, no exchange during the product" % (n, world),
                           "nnz_in": int(nnzb), "nnz_out": int(nnzc), "nnz_out_closed_form": expect,
                           "structure_count_matches_closed_form": (int(nnzc) == expect) if expect else None},
                "clocks": clocks, "gpu_launches": int(launches),
                "roofline": {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak,
                             "traffic": None, "peak_source": peak_src, "kernel": "k_mm_merge (numeric)",
                             "note": "per-GPU share of the one-pass byte floor (B counted once per rank)"}}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def bench_powerlaw_sharded(args, rank, world, local_rank, dist):
    """cfg 4 over `world` GPUs: R-MAT rows cut into blocks of equal nonzeros, x sliced by the same
    bounds; every product all-gathers x over NVLink (k_allgather_pull) and runs the balanced kernel."""
    import torch
    import sparse_linear_b200 as sl
    from . import generators as g
    import bench as B
    scale = args.grid or 24
    stream = torch.cuda.current_stream()
    sl.set_stream(stream.cuda_stream)
    cpu_group = dist.new_group(backend="gloo") if world > 1 else None
    n, r, c, v = g.rmat_triplets(scale, 16, seed=0)          # every rank generates the same edge list
    splits = balanced_splits(np.bincount(r, minlength=n), world)
    r0, r1 = int(splits[rank]), int(splits[rank + 1])
    keep = (r >= r0) & (r < r1)
    rl, cl, vl = (r[keep] - r0).astype(np.int32), c[keep], v[keep]
    del r, c, v, keep
    local = sl.compress(r1 - r0, n, rl, cl, vl)               # device assembly, duplicates summed
    nnz_local = sl.nonZero(local)
    sh = ShardedSpMV(local, splits, rank, world, dist, cpu_group, mode="allgather")
    xfull = g.uniform_vector(n, seed=0)
    sh.set_x(xfull[r0:r1])
    for _ in range(max(args.warmup, 3)):
        sh.mulv()
    torch.cuda.synchronize()
    want = np.bincount(rl, weights=vl * xfull[cl], minlength=r1 - r0)
    got = sh.y()
    err = float(np.linalg.norm(got - want) / max(np.linalg.norm(want), 1e-300))
    del rl, cl, vl
    sampler = B.ClockSampler(local_rank)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    l0 = sl.launch_count()
    sampler.start()
    e0.record(stream)
    for _ in range(args.steps):
        sh.mulv()
    e1.record(stream)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    clocks = sampler.stop()
    launches = sl.launch_count() - l0
    ms = _max_over_ranks(e0.elapsed_time(e1) / args.steps, dist, world)
    nnz, errsq = _sum_over_ranks([nnz_local, err * err], dist, world)
    abytes = g.spmv_bytes(n, n, int(nnz))
    value = abytes / (ms * 1e-3) / 1e9
    if rank == 0:
        peak, peak_src = B.peaks()
        per_gpu = value / world
        line = {"metric": B.METRIC, "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic (seeded numpy PCG64)",
                "config": {"workload": "cfg4: R-MAT scale %d edge factor 16, Double SpMV, rows in %d blocks of equal "
                                       "nonzeros; x all-gathered over NVLink peer memory every product (%.0f MB inbound "
                                       "per GPU), then the balanced kernel" % (scale, world, n * 8 * (world - 1) / world / 1e6),
                           "rows": n, "nnz": int(nnz), "algorithmic_bytes_per_step": abytes,
                           "plan": local.device().plan_info(), "splits": [int(x) for x in splits],
                           "rel_err_vs_host_sum_rms_over_ranks": float(np.sqrt(errsq / world))},
                "clocks": clocks, "gpu_launches": int(launches), "e2e": None,
                "roofline": {"bound": "hbm", "achieved": per_gpu, "peak": peak, "unit": "GB/s", "frac": per_gpu / peak,
                             "traffic": None, "peak_source": peak_src,
                             "kernel": "k_allgather_pull + k_spmv_bal",
                             "note": "per-GPU figure; the step includes the barrier and the all-gather of x",
                             "barrier_timed_out": sh.timed_out()},
                "cpu_baseline": None}
        print(json.dumps(line), flush=True)
    sh.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def bench_structural_sharded(args, rank, world, local_rank, dist):
    """cfg 3 transpose / assembly over `world` GPUs: column blocks in, column blocks out, one
    all-to-all-v (NCCL), local kernels on each GPU.  The pieces are cut, shipped and stacked through
    host memory this round, so the figure is the whole step (device kernels + PCIe + host stacking),
    wall clock, max over ranks."""
    import torch
    import sparse_linear_b200 as sl
    from . import generators as g
    import bench as B
    n = args.grid or 128
    a = g.laplace3d_7pt(n)
    sl.set_stream(torch.cuda.current_stream().cuda_stream)
    cs = uniform_splits(n, world) * (n * n)
    c0, c1 = int(cs[rank]), int(cs[rank + 1])
    blk = col_block(a, c0, c1)
    steps = max(1, min(args.steps, 5))
    if args.workload == "cfg3t":
        def step():
            return sharded_transpose(blk, cs, dist)
        ref = blk          # the Laplacian is symmetric: block c of the transpose is block c of A
        units, unit, metric = int(a.colptr[-1]), "nnz/s", "transpose_nnz_per_s"
    else:
        r, c, v = g.split_duplicates(a, seed=0)       # every rank generates the same shuffled triplets
        lo, hi = (len(r) * rank) // world, (len(r) * (rank + 1)) // world
        r, c, v = r[lo:hi], c[lo:hi], v[lo:hi]

        def step():
            m = sharded_compress(a.nrows, r, c, v, cs, rank, dist)
            return CSC(m.nrows, m.ncols, m.pointers, m.indices, m.values)
        ref = blk          # 0.25 v + 0.75 v == v exactly
        units, unit, metric = 2 * int(a.colptr[-1]), "triplets/s", "assembly_triplets_per_s"
    got = step()
    ok = bool(np.array_equal(got.colptr, ref.colptr) and np.array_equal(got.rowidx, ref.rowidx)
              and np.array_equal(got.values, ref.values))
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    l0 = sl.launch_count()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    torch.cuda.synchronize()
    ms = _max_over_ranks((time.perf_counter() - t0) / steps * 1e3, dist, world)
    oks = _sum_over_ranks([1.0 if ok else 0.0], dist, world)[0]
    if rank == 0:
        line = {"metric": metric, "value": units / (ms * 1e-3), "unit": unit, "n_gpus": world, "steps": steps, "warmup": 1,
                "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "u32 keys + f64 values", "data": "synthetic (seeded numpy PCG64)",
                "config": {"workload": "cfg3: %s of the 3D 7-point Laplacian %d^3 held as %d column blocks; local %s on each GPU, "
                                       "one all-to-all-v over NCCL, pieces cut and stacked on the host"
                                       % ("transpose" if args.workload == "cfg3t" else "triplet assembly (2x duplicates)", n, world,
                                          "slc_d_transpose" if args.workload == "cfg3t" else "slc_d_compress"),
                           "blocks_equal_expected_bit_exact_on_ranks": int(oks), "timing": "wall clock, max over ranks"},
                "gpu_launches": int(sl.launch_count() - l0), "roofline": None}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
