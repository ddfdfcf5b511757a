"""bench.py --gpus N (N > 1): one process per GPU (torchrun), the row-partitioned products of
sparse-linear_b200/sharded.py.  Rank 0 prints ONE JSON line: the headline is cfg 5 SpMV strong-scaled
over z-slabs (or weak with --scaling weak); `suite` adds cfg 5 SpMM, cfg 3 SpGEMM, cfg 4 SpMV and the
sharded transpose / assembly at the same N.  Every entry carries an oracle verdict computed AFTER its
timed loop on what the loop left in HBM, AND-reduced over the ranks.

Timing protocol: barrier + synchronize, then -- untimed, on the device -- a rendezvous of all GPUs
(slc_peer_barrier_all: every rank waits for every rank in a one-warp kernel over peer memory) and ONE
aligning product, event, K products, event, synchronize, barrier; device time, max over ranks.  Without
the device-side rendezvous the first timed steps absorb the skew with which the host processes leave the
NCCL barrier (hundreds of microseconds against a 45 us step at 8 GPUs; the products themselves only wait
for their neighbours, so a late rank holds its neighbours' first event back but not the far ranks').
By default the K products are captured once in a CUDA graph, the graph is replayed once untimed (which
uploads it) and launched inside the timed region (--no-graph: K host launches)."""
import ctypes as C
import json
import os
import time
import traceback

import numpy as np

import bench as B
from sparse_linear_b200 import _capi
from sparse_linear_b200 import generators as g
from sparse_linear_b200 import sharded as S
from sparse_linear_b200.generators import CSC
from sparse_linear_b200.matrix import Matrix, _check


def _reduce(vals, dist, world, op="sum"):
    import torch
    t = torch.tensor([float(v) for v in vals], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op={"sum": dist.ReduceOp.SUM, "max": dist.ReduceOp.MAX, "min": dist.ReduceOp.MIN}[op])
    return [float(v) for v in t.tolist()]


def timed_region(steps, step, align, stream, dist, world, local_rank, sl, use_graph, rendezvous=None):
    """The protocol of the module docstring.  Returns (ms per step max over ranks, clocks, launches, how)."""
    import torch
    how = "host launches"
    graph = None
    if use_graph:
        try:
            cap = torch.cuda.Stream()
            sl.set_stream(cap.cuda_stream)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.stream(cap):
                cap.synchronize()
                with torch.cuda.graph(graph, stream=cap):
                    for _ in range(steps):
                        step()
            how = "one CUDA graph of %d products" % steps
        except Exception as ex:  # noqa: BLE001
            graph = None
            how = "host launches (graph capture failed: %r)" % (ex,)
        finally:
            sl.set_stream(stream.cuda_stream)
    if graph is not None:
        graph.replay()            # untimed: the first launch uploads the graph to the device
        torch.cuda.synchronize()
    sampler = B.ClockSampler(local_rank)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler.start()
    if rendezvous is not None and world > 1:
        rendezvous()              # untimed, on the device: every GPU waits for every GPU
    la = sl.launch_count()
    align()                       # untimed: one product, so the boundary waits of the first timed one are warm
    l0 = sl.launch_count()
    per_step = l0 - la
    e0.record(stream)
    if graph is not None:
        graph.replay()
    else:
        for _ in range(steps):
            step()
    e1.record(stream)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    clocks = sampler.stop()
    launches = (sl.launch_count() - l0) if graph is None else steps * max(1, per_step)
    mine = e0.elapsed_time(e1) / steps
    ms = _reduce([mine], dist, world, "max")[0]
    rank = dist.get_rank() if world > 1 else 0
    per_rank = _reduce([mine if r == rank else 0.0 for r in range(world)], dist, world, "sum")
    how += "; ms per step by rank: " + " ".join("%.4f" % v for v in per_rank)
    return ms, clocks, int(launches), how


def row_block_of_symmetric(colblock):
    """A[r0:r1, :] of a symmetric A given A[:, r0:r1]: the oracle's transpose (values bit for bit)."""
    import oracle
    return oracle.transpose(oracle.csc(*colblock))


# ---- cfg 5: SpMV / SpMM over z-slabs --------------------------------------------------------------------

def stencil_entry(args, rank, world, local_rank, dist, sl, stream, cpu_group, k):
    import torch
    import oracle
    n = args.grid or 200
    shape = (n * world, n, n) if args.scaling == "weak" else (n, n, n)
    plane = shape[1] * shape[2]
    ntot = shape[0] * plane
    splits = S.uniform_splits(shape[0], world) * plane        # z-slabs
    r0, r1 = int(splits[rank]), int(splits[rank + 1])
    # A is symmetric, so the row block A[r0:r1, :] is the transpose of the column block A[:, r0:r1]
    colblock = g.stencil3d_27pt(shape, col_range=(r0, r1))
    cb = Matrix(colblock.ncols, colblock.nrows, colblock.colptr, colblock.rowidx, colblock.values)
    local = sl.transpose(cb)                                 # (r1-r0) x ntot, on the device
    del cb
    sh = S.ShardedSpMV(local, splits, rank, world, dist, cpu_group, k=k)
    xfull = g.uniform_vector(ntot, seed=0)
    nnz_local = int(colblock.colptr[-1])
    if k is None:
        xloc = np.ascontiguousarray(xfull[r0:r1])
        abytes_local = g.spmv_bytes(r1 - r0, r1 - r0, nnz_local)   # every x entry is read from HBM once, by its owner
    else:
        # column j of the block is x scaled by 2^-(j % 8) * (1 + j // 8): exact scalings of one vector, so
        # the oracle's product of a few columns checks them all the same way
        colscale = (2.0 ** -(np.arange(k) % 8)) * (1 + np.arange(k) // 8)
        xloc = np.asfortranarray(xfull[r0:r1, None] * colscale[None, :])
        abytes_local = g.spmm_bytes(r1 - r0, r1 - r0, nnz_local, k)
    sh.set_x(xloc)
    sh.cur ^= 1
    sh.set_x(xloc)               # both operand buffers hold x: the timed products alternate between them
    sh.cur ^= 1
    abytes, nnz = _reduce([abytes_local, nnz_local], dist, world)
    steps = args.steps if k is None else max(3, min(args.steps, 20))
    for _ in range(max(args.warmup, 3)):
        sh.mulv()
    torch.cuda.synchronize()
    sh.stats(reset=True)
    ms, clocks, launches, how = timed_region(steps, sh.mulv, sh.mulv, stream, dist, world, local_rank, sl,
                                             use_graph=not args.no_graph and not sh.use_nccl_barrier,
                                             rendezvous=None if sh.use_nccl_barrier else sh.barrier_all)
    st = sh.stats()
    value = abytes / (ms * 1e-3) / 1e9
    # parity AFTER the timed loop, on what it left in HBM: this rank's rows against the oracle
    got = sh.y()
    oblk = row_block_of_symmetric(colblock)
    if k is None:
        want = oracle.mulv(oblk, xfull)
        ok = bool(np.array_equal(got, want))
    else:
        cols = [0, k // 3, (2 * k) // 3, k - 1]
        Xc = np.asfortranarray(xfull[:, None] * colscale[None, cols])
        want = oracle.mulm(oblk, Xc)
        ok = bool(np.array_equal(got[:, cols], want))
    all_ok = _reduce([1.0 if ok else 0.0], dist, world, "min")[0] == 1.0
    waits = _reduce([st[2] / max(st[1], 1), st[3]], dist, world, "max")

    e2e = None
    if not args.no_e2e:
        lib = _capi.lib()
        ncol = 1 if k is None else k
        xp = torch.from_numpy(np.ascontiguousarray(xloc.T if k is not None else xloc)).pin_memory()
        yp = torch.empty((r1 - r0) * ncol, dtype=torch.float64).pin_memory()
        steps_e2e = max(3, min(args.steps, 20 if k is None else 5))

        def hstep():
            # the operand goes into the buffer the previous product did not read (double buffering: no peer
            # can still be reading it), then the product, then the result comes back
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
        dt = _reduce([(time.perf_counter() - t0) / steps_e2e], dist, world, "max")[0]
        got2 = yp.numpy() if k is None else yp.numpy().reshape(k, r1 - r0).T
        ok2 = bool(np.array_equal(got2, want) if k is None else np.array_equal(got2[:, cols], want))
        all_ok2 = _reduce([1.0 if ok2 else 0.0], dist, world, "min")[0] == 1.0
        e2e = {"value": abytes / dt / 1e9, "unit": "GB/s", "ms_per_step": dt * 1e3, "steps": steps_e2e,
               "h2d_bytes_per_step": int(ntot * 8 * ncol), "d2h_bytes_per_step": int(ntot * 8 * ncol),
               "api": "ShardedSpMV (slc_memcpy_h2d + slc_d_%s_dist_dev + slc_memcpy_d2h per rank)" % ("mulv" if k is None else "mulm"),
               "matches_oracle_bit_exact": all_ok2}
    per_gpu = value / world
    entry = {
        "name": "cfg5_spmv" if k is None else "cfg5_spmm",
        "metric": B.METRIC if k is None else "spmm_effective_hbm_gbps", "value": value, "unit": "GB/s", "ms_per_step": ms,
        "steps": steps, "dtype": "f64",
        "config": {"workload": ("cfg5: 3D 27-point stencil %dx%dx%d, Double SpMV, row-partitioned in z-slabs over %d GPU(s); "
                                "halo entries gathered from peer HBM over NVLink inside the product kernel; signal and "
                                "neighbour-only waits (peer-memory flags) fused into the same launch" % (shape + (world,)))
                   if k is None else
                   ("cfg5: 3D 27-point stencil %dx%dx%d, Double SpMM k=64, row-partitioned in z-slabs over %d GPU(s); "
                    "signal, interior rows, wait, boundary rows gathering the peers' rows of X over NVLink" % (shape + (world,))),
                   "rows": ntot, "nnz": int(nnz), "algorithmic_bytes_per_step": abytes, "plan": local.device().plan_info(),
                   "l2": "inputs larger than L2 (%.0f MB per GPU per step)" % (abytes / world / 1e6),
                   "pct_of_nominal_8TBs_per_gpu": 100.0 * per_gpu / B.NOMINAL_HBM_GBS,
                   "timed_region": how + "; untimed before the first event: a device-side rendezvous of all ranks and one aligning product",
                   "operand": "two exported buffers per rank, alternated per product (double buffering)",
                   "dependency_mask_rank0": sh.deps},
        "clocks": clocks, "gpu_launches": launches, "e2e": e2e,
        "parity": {"oracle": "oracle.%s on every rank's row block, after the timed loop" % ("mulv" if k is None else "mulm (4 columns)"),
                   "bit_exact_all_ranks": all_ok, "ok": all_ok, "required": "bit-exact"},
        "roofline": B.roofline(per_gpu, "k_spmv_sell_dist" if k is None else "k_spmm_panel/brick + k_spmm_boundary",
                               note="per-GPU figure; the step includes signal and waits",
                               barrier="nccl all-reduce" if sh.use_nccl_barrier else "peer-memory flags, neighbour-only waits",
                               barrier_timed_out=sh.timed_out(),
                               boundary_wait_us={"mean_per_block_max_over_ranks": waits[0] / 1e3, "longest_max_over_ranks": waits[1] / 1e3}),
        "cpu_baseline": None,
    }
    sh.close()
    del local
    return entry


# ---- cfg 3: SpGEMM over ranks: column blocks of B (default) or row blocks of A ------------------------------

def spgemm_entry(args, rank, world, local_rank, dist, sl, stream):
    """C = A A on the cfg 3 Laplacian.  SLC_MM_PARTITION=cols (default): rank p computes C[:, J_p] = A B[:, J_p]
    with A resident everywhere -- every per-column pass of the product shrinks with the block and the blocks
    concatenate as they are.  =rows: C[I_p, :] = A[I_p, :] B with B replicated (round 1: the per-column passes
    over all of B's columns stay on every rank: 2.8 x at 8 GPUs)."""
    import oracle
    n = args.grid or 128
    a = g.laplace3d_7pt(n)
    by_cols = os.environ.get("SLC_MM_PARTITION", "cols") != "rows"
    splits = S.uniform_splits(n, world) * (n * n)           # z-slabs: equal work per rank
    lo, hi = int(splits[rank]), int(splits[rank + 1])
    whole = Matrix(a.ncols, a.nrows, a.colptr, a.rowidx, a.values)
    blk = S.col_block(a, lo, hi) if by_cols else S.row_block(a, lo, hi)
    local = Matrix(blk.ncols, blk.nrows, blk.colptr, blk.rowidx, blk.values)
    keep = {}

    def step():
        keep["c"] = sl.mm(whole, local) if by_cols else sl.mm(local, whole)
    steps = max(3, min(args.steps, 20))
    for _ in range(3):
        step()
    ms, clocks, launches, how = timed_region(steps, step, step, stream, dist, world, local_rank, sl, use_graph=False)
    c = keep["c"]
    nnzc_local = sl.nonZero(c)
    want = oracle.mm(oracle.csc(*a), oracle.csc(*blk)) if by_cols else oracle.mm(oracle.csc(*blk), oracle.csc(*a))
    ok = B.same_matrix(c, want)
    all_ok = _reduce([1.0 if ok else 0.0], dist, world, "min")[0] == 1.0
    nnzc, nnzblk = _reduce([nnzc_local, blk.colptr[-1]], dist, world)
    nnzfull = int(a.colptr[-1])
    floor_bytes = (nnzblk + nnzfull * world + nnzc) * 12 + 3 * (a.ncols + 1) * 4 * world
    gbs = floor_bytes / (ms * 1e-3) / 1e9 / world
    how_cut = ("B in %d column blocks (z-slabs), A resident on every rank" if by_cols else
               "A in %d row blocks (z-slabs), B replicated") % world
    return {"name": "cfg3_spgemm", "metric": "spgemm_output_nnz_per_s", "value": nnzc / (ms * 1e-3), "unit": "nnz/s",
            "ms_per_step": ms, "steps": steps, "dtype": "f64", "clocks": clocks, "gpu_launches": launches,
            "config": {"workload": "cfg3: SpGEMM A*A, 3D 7-point Laplacian %d^3, %s, no exchange during the product" % (n, how_cut),
                       "partition": "cols" if by_cols else "rows", "nnz_in": nnzfull, "nnz_out": int(nnzc)},
            "parity": {"oracle": "oracle.mm on every rank's block (%s), after the timed loop" %
                                 ("A, B[:, J]" if by_cols else "A[I, :], B"),
                       "bit_exact_all_ranks": all_ok, "ok": all_ok, "required": "structure and values bit-exact"},
            "roofline": B.roofline(gbs, "k_mm_merge (numeric)", note="per-GPU share of the one-pass byte floor (the resident operand counted once per rank)")}


# ---- cfg 4: power-law SpMV, x all-gathered over peer memory ----------------------------------------------

def powerlaw_entry(args, rank, world, local_rank, dist, sl, stream, cpu_group):
    import torch
    import oracle
    scale = args.grid or 24
    n, r, c, v = B.rmat_device(scale)                         # every rank generates the same edge list on its GPU
    counts = torch.bincount(r, minlength=n).cpu().numpy()
    splits = S.balanced_splits(counts, world)
    r0, r1 = int(splits[rank]), int(splits[rank + 1])
    keep = (r >= r0) & (r < r1)
    rl, cl, vl = (r[keep] - r0).contiguous(), c[keep].contiguous(), v[keep].contiguous()
    del r, c, v, keep
    local = sl.compress_dev(r1 - r0, n, rl.numel(), rl.data_ptr(), cl.data_ptr(), vl.data_ptr(), np.float64)
    sl.synchronize()
    del rl, cl, vl
    torch.cuda.empty_cache()
    nnz_local = sl.nonZero(local)
    # Who owns which part of x is independent of who owns which rows.  With x cut at the row bounds the rank that
    # owns the sparsest rows owns most of x (58 of 134 MB at 8 ranks) and the other seven all pull that slice from
    # it at once: 408 MB through one GPU's NVLink egress, 0.45 ms at line rate -- more than the product itself.
    # Equal slices of x (SLC_AG_XSPLIT=uniform, default) make every rank serve 7 x 16.8 MB instead.  =rows keeps
    # the x partition equal to the row partition (what an iteration y -> x without redistribution wants).
    x_uniform = os.environ.get("SLC_AG_XSPLIT", "uniform") != "rows"
    xs = S.uniform_splits(n, world) if x_uniform else splits
    sh = S.ShardedSpMV(local, xs, rank, world, dist, cpu_group, mode="allgather")
    xfull = g.uniform_vector(n, seed=0)
    xloc = np.ascontiguousarray(xfull[int(xs[rank]):int(xs[rank + 1])])
    sh.set_x(xloc)
    sh.cur ^= 1
    sh.set_x(xloc)
    sh.cur ^= 1
    for _ in range(max(args.warmup, 3)):
        sh.mulv()
    torch.cuda.synchronize()
    steps = max(3, min(args.steps, 50))
    ms, clocks, launches, how = timed_region(steps, sh.mulv, sh.mulv, stream, dist, world, local_rank, sl, use_graph=False,
                                             rendezvous=None if sh.use_nccl_barrier else sh.barrier_all)
    got = sh.y()
    oblk = oracle.csc(local.nrows, local.ncols, local.pointers, local.indices, local.values)
    want = oracle.mulv(oblk, xfull)
    err = float(np.linalg.norm(got - want) / max(np.linalg.norm(want), 1e-300))
    worst = _reduce([err], dist, world, "max")[0]
    nnz = _reduce([nnz_local], dist, world)[0]
    abytes = g.spmv_bytes(n, n, int(nnz))
    value = abytes / (ms * 1e-3) / 1e9
    per_gpu = value / world
    entry = {"name": "cfg4_spmv", "metric": B.METRIC, "value": value, "unit": "GB/s", "ms_per_step": ms, "steps": steps,
             "dtype": "f64", "clocks": clocks, "gpu_launches": launches,
             "config": {"workload": "cfg4: R-MAT scale %d edge factor 16, Double SpMV, rows in %d blocks of equal nonzeros, x owned in %s; x "
                                    "all-gathered over NVLink peer memory every product (%.0f MB inbound per GPU), then the "
                                    "balanced kernel" % (scale, world, "equal slices" if x_uniform else "the same blocks",
                                                         n * 8 * (world - 1) / world / 1e6),
                        "rows": n, "nnz": int(nnz), "algorithmic_bytes_per_step": abytes, "plan": local.device().plan_info(),
                        "splits": [int(x) for x in splits], "x_splits": [int(x) for x in xs]},
             "parity": {"oracle": "oracle.mulv on every rank's device-assembled row block, after the timed loop",
                        "max_rel_err_over_ranks": worst, "ok": worst <= 1e-12, "required": "1e-12"},
             "roofline": B.roofline(per_gpu, "k_allgather_pull + k_spmv_bal",
                                    note="per-GPU figure; the step includes the barrier and the all-gather of x",
                                    barrier_timed_out=sh.timed_out())}
    sh.close()
    return entry


# ---- cfg 3: transpose and assembly over ranks, on the device ---------------------------------------------

def structural_entry(args, rank, world, local_rank, dist, sl, stream, what):
    import torch
    import oracle
    n = args.grid or 128
    a = g.laplace3d_7pt(n)
    cs = S.uniform_splits(n, world) * (n * n)
    c0, c1 = int(cs[rank]), int(cs[rank + 1])
    blk = S.col_block(a, c0, c1)
    steps = max(2, min(args.steps, 5))
    if what == "transpose":
        local = Matrix(blk.ncols, blk.nrows, blk.colptr, blk.rowidx, blk.values)
        local.device()

        def step():
            return S.sharded_transpose_device(local, cs, dist)
        ref = blk          # the Laplacian is symmetric: block c of the transpose is block c of A
        units, unit, metric = int(a.colptr[-1]), "nnz/s", "transpose_nnz_per_s"
    else:
        r, c, v = g.split_duplicates(a, seed=0)       # every rank generates the same shuffled triplets
        lo, hi = (len(r) * rank) // world, (len(r) * (rank + 1)) // world
        dr = torch.from_numpy(r[lo:hi].astype(np.int32)).cuda()
        dc = torch.from_numpy(c[lo:hi].astype(np.int32)).cuda()
        dv = torch.from_numpy(v[lo:hi]).cuda()

        def step():
            return S.sharded_compress_device(a.nrows, dr, dc, dv, cs, rank, dist)
        ref = blk          # 0.25 v + 0.75 v == v exactly
        units, unit, metric = 2 * int(a.colptr[-1]), "triplets/s", "assembly_triplets_per_s"
    got = step()
    torch.cuda.synchronize()
    ok = bool(np.array_equal(got.pointers, ref.colptr) and np.array_equal(got.indices, ref.rowidx)
              and np.array_equal(got.values, ref.values))
    keep = {}

    def tstep():
        keep["m"] = step()
    ms, clocks, launches, how = timed_region(steps, tstep, tstep, stream, dist, world, local_rank, sl, use_graph=False)
    all_ok = _reduce([1.0 if ok else 0.0], dist, world, "min")[0] == 1.0
    return {"name": "cfg3_" + what, "metric": metric, "value": units / (ms * 1e-3), "unit": unit, "ms_per_step": ms, "steps": steps,
            "dtype": "u32 keys + f64 values", "clocks": clocks, "gpu_launches": launches,
            "config": {"workload": "cfg3: %s of the 3D 7-point Laplacian %d^3 held as %d column blocks: local device kernels, pieces cut "
                                   "on the device, ONE all-to-all-v of device buffers (NCCL), device stacking / assembly"
                                   % ("transpose" if what == "transpose" else "triplet assembly (2x duplicates)", n, world)},
            "parity": {"oracle": "the known answer (this rank's column block of A, bit for bit)", "bit_exact_all_ranks": all_ok,
                       "ok": all_ok, "required": "bit-exact"},
            "roofline": None}


def main(args, rank, world, local_rank, dist):
    import torch
    import sparse_linear_b200 as sl
    stream = torch.cuda.current_stream()
    sl.set_stream(stream.cuda_stream)
    cpu_group = dist.new_group(backend="gloo") if world > 1 else None
    w = args.workload

    def run(fn, name):
        t0 = time.perf_counter()
        try:
            e = fn()
        except Exception as ex:  # noqa: BLE001
            e = {"name": name, "error": repr(ex), "trace": traceback.format_exc()[-1500:]}
            try:
                torch.cuda.synchronize()
            except Exception:  # noqa: BLE001
                pass
        e["wall_s"] = round(time.perf_counter() - t0, 2)
        torch.cuda.empty_cache()
        return e

    table = {
        "cfg5": lambda: stencil_entry(args, rank, world, local_rank, dist, sl, stream, cpu_group, None),
        "cfg5mm": lambda: stencil_entry(args, rank, world, local_rank, dist, sl, stream, cpu_group, 64),
        "cfg3": lambda: spgemm_entry(args, rank, world, local_rank, dist, sl, stream),
        "cfg4": lambda: powerlaw_entry(args, rank, world, local_rank, dist, sl, stream, cpu_group),
        "cfg3t": lambda: structural_entry(args, rank, world, local_rank, dist, sl, stream, "transpose"),
        "cfg3a": lambda: structural_entry(args, rank, world, local_rank, dist, sl, stream, "assembly"),
    }
    if w not in table:
        raise SystemExit("bench.py --gpus N: workload %s has no multi-GPU form" % w)
    head = run(table[w], w)
    suite = []
    if w == "cfg5" and args.suite == "auto" and not args.grid and args.scaling == "strong":
        for name in ("cfg5mm", "cfg3", "cfg4", "cfg3t", "cfg3a"):
            suite.append(run(table[name], name))
    if rank == 0:
        line = {"metric": head.get("metric", B.METRIC), "value": head.get("value"), "unit": head.get("unit", "GB/s"), "n_gpus": world,
                "steps": head.get("steps", args.steps), "warmup": max(args.warmup, 3), "ms_per_step": head.get("ms_per_step"),
                "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": head.get("dtype", "f64"),
                "data": "synthetic (seeded numpy PCG64 / torch Philox)"}
        for k in ("config", "clocks", "gpu_launches", "e2e", "parity", "roofline", "cpu_baseline", "error", "trace"):
            if k in head:
                line[k] = head[k]
        if suite:
            line["suite"] = suite
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
