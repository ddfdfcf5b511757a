"""Worker of tests/test_gpu_multiproc.py: one process per rank (gloo for the object exchange), CUDA IPC
mappings of the peers' operand buffers, the fused signal / wait inside the product kernel, and x
REWRITTEN before every product -- the path bench.py --gpus N takes, under the oracle.

    python mp_sharded_worker.py <rank> <world> <port> <mode> <steps> <out.json>

mode: spmv | spmm | allgather.  Every product of every step is compared with the oracle on this
rank's row block (bit-exact where one thread owns a row, 1e-12 for the balanced kernel).
"""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import numpy as np  # noqa: E402


def main():
    rank, world, port, mode, steps, out = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4], int(sys.argv[5]), sys.argv[6]
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import oracle
    import sparse_linear_b200 as sl
    from sparse_linear_b200 import generators as g
    from sparse_linear_b200 import sharded
    from helpers import rmat_csc
    lib = sl._capi.lib()
    ndev = sl.device_count()
    sl.set_device(rank % ndev)
    result = {"rank": rank, "device": rank % ndev, "ok": False}
    if mode in ("transpose", "assembly"):
        structural(rank, world, mode, out, dist, sl, g, sharded, oracle)
        return
    try:
        if mode == "allgather":
            gm = rmat_csc(12, 8, seed=5)
            rs = sharded.balanced_splits(np.bincount(gm.rowidx, minlength=gm.nrows), world)
            k = None
        else:
            shape = (4 * world, 16, 64)                      # z-slabs of 4 planes; lines are whole slices
            gm = g.stencil3d_27pt(shape)
            rs = sharded.uniform_splits(shape[0], world) * (16 * 64)
            k = 6 if mode == "spmm" else None
        r0, r1 = int(rs[rank]), int(rs[rank + 1])
        blk = sharded.row_block(gm, r0, r1)
        oblk = oracle.csc(*blk)
        local = sl.Matrix(blk.ncols, blk.nrows, blk.colptr, blk.rowidx, blk.values)
        if mode != "allgather":
            local.device().set_spmv_kernel(1)
        sh = sharded.ShardedSpMV(local, rs, rank, world, dist, None, k=k, mode="allgather" if mode == "allgather" else "halo")
        result["deps"] = sh.deps
        result["plan"] = local.device().plan_info()
        ncol = 1 if k is None else k
        nloc, nrow = r1 - r0, blk.nrows
        # every step's operand, generated identically on all ranks; all slices go to the device up front so
        # the timed loop is device-side only (no host synchronisation between the products of a burst)
        rng = np.random.Generator(np.random.PCG64(1234))
        xs = rng.uniform(-1, 1, (steps, ncol, gm.ncols))
        mine = np.ascontiguousarray(xs[:, :, r0:r1])
        dxs, dys = C.c_void_p(), C.c_void_p()
        assert lib.slc_dev_alloc(C.byref(dxs), mine.nbytes) == 0
        assert lib.slc_dev_alloc(C.byref(dys), steps * ncol * nrow * 8) == 0
        assert lib.slc_memcpy_h2d(dxs, C.c_void_p(mine.ctypes.data), mine.nbytes) == 0
        dist.barrier()
        for s in range(steps):
            # rewrite the operand buffer the next product reads (device-to-device, asynchronous), multiply,
            # keep the result: nothing here waits for the GPU, so a rank may run ahead of its peers as far
            # as the protocol lets it
            assert lib.slc_memcpy_d2d(sh.x_ptr, C.c_void_p(dxs.value + s * ncol * nloc * 8), ncol * nloc * 8) == 0
            sh.mulv()
            assert lib.slc_memcpy_d2d(C.c_void_p(dys.value + s * ncol * nrow * 8), sh.y_ptr, ncol * nrow * 8) == 0
        sl.synchronize()
        got = np.empty((steps, ncol, nrow))
        assert lib.slc_memcpy_d2h(C.c_void_p(got.ctypes.data), dys, got.nbytes) == 0
        bad = []
        for s in range(steps):
            for j in range(ncol):
                want = oracle.mulv(oblk, np.ascontiguousarray(xs[s, j]))
                if mode == "allgather":
                    good = np.linalg.norm(got[s, j] - want) <= 1e-12 * np.linalg.norm(want)
                else:
                    good = np.array_equal(got[s, j], want)
                if not good:
                    bad.append((s, j))
        result["bad"] = bad[:10]
        result["timed_out"] = sh.timed_out()
        result["stats"] = sh.stats()
        result["ok"] = not bad and not result["timed_out"]
        dist.barrier()
        sh.close()
    except Exception as e:  # noqa: BLE001
        result["error"] = repr(e)
    json.dump(result, open(out, "w"))
    try:
        dist.destroy_process_group()
    except Exception:
        pass


def structural(rank, world, mode, out, dist_gloo, sl, g, sharded, oracle):
    """Sharded transpose / assembly with every entry on the GPUs: NCCL all-to-all-v of device buffers
    (needs one GPU per rank), blocks compared with the oracle's."""
    import torch
    import torch.distributed as dist
    result = {"rank": rank, "ok": False}
    try:
        torch.cuda.set_device(rank)
        nccl = dist.new_group(backend="nccl")
        rng = np.random.default_rng(77)
        from helpers import random_csc
        m = random_csc(rng, 500, 420, 0.05)              # every rank builds the same matrix
        cs = sharded.uniform_splits(m.ncols, world)
        rs = sharded.uniform_splits(m.nrows, world)
        if mode == "transpose":
            blk = sharded.col_block(m, cs[rank], cs[rank + 1])
            local = sl.Matrix(blk.ncols, blk.nrows, blk.colptr, blk.rowidx, blk.values)
            got = sharded.sharded_transpose_device(local, rs, dist, nccl, col_splits=cs)
            ref = sharded.col_block(oracle.transpose(oracle.csc(*m)), rs[rank], rs[rank + 1])
        else:
            n = 20000
            r = rng.integers(0, m.nrows, n).astype(np.int32)
            c = rng.integers(0, m.ncols, n).astype(np.int32)
            v = rng.integers(-5, 6, n).astype(np.float64)
            lo, hi = n * rank // world, n * (rank + 1) // world
            got = sharded.sharded_compress_device(m.nrows, torch.from_numpy(r[lo:hi]).cuda(), torch.from_numpy(c[lo:hi]).cuda(),
                                                  torch.from_numpy(v[lo:hi]).cuda(), cs, rank, dist, nccl)
            ref = sharded.col_block(oracle.compress(m.nrows, m.ncols, r, c, v), cs[rank], cs[rank + 1])
        result["ok"] = bool(np.array_equal(got.pointers, ref.colptr) and np.array_equal(got.indices, ref.rowidx)
                            and np.array_equal(got.values, ref.values))
        dist.barrier(group=nccl)
    except Exception as e:  # noqa: BLE001
        import traceback
        result["error"] = repr(e) + traceback.format_exc()[-800:]
    json.dump(result, open(out, "w"))
    try:
        dist.destroy_process_group()
    except Exception:
        pass


if __name__ == "__main__":
    main()
