"""Host-side logic of the row-partitioned SpMV (partitioning, row blocks, handle exchange), run as
world_size-2 gloo processes on the CPU.  The device kernels are covered by tests/test_gpu_sharded.py."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
from helpers import random_csc
from sparse_linear_b200 import generators as g
from sparse_linear_b200 import sharded


def test_uniform_and_balanced_splits():
    assert sharded.uniform_splits(200, 8).tolist() == [0, 25, 50, 75, 100, 125, 150, 175, 200]
    assert sharded.uniform_splits(10, 3).tolist() == [0, 3, 6, 10]
    assert sharded.uniform_splits(100, 3, align=32).tolist() == [0, 32, 64, 100]
    s = sharded.uniform_splits(5, 8)
    assert s[0] == 0 and s[-1] == 5 and np.all(np.diff(s) >= 0)
    nnz = np.array([1, 1, 1, 1, 100, 1, 1, 1, 1, 1])
    b = sharded.balanced_splits(nnz, 2)
    assert b[0] == 0 and b[-1] == 10 and 4 <= b[1] <= 5
    b4 = sharded.balanced_splits(np.ones(16), 4)
    assert b4.tolist() == [0, 4, 8, 12, 16]
    assert np.all(np.diff(sharded.balanced_splits(np.zeros(7), 3)) >= 0)


def test_row_block_partitions_the_product():
    rng = np.random.default_rng(0)
    m = random_csc(rng, 50, 40, 0.2)
    x = rng.uniform(-1, 1, 40)
    full = oracle.mulv(m, x)
    splits = sharded.uniform_splits(50, 3)
    parts = []
    for p in range(3):
        blk = sharded.row_block(m, splits[p], splits[p + 1])
        assert blk.nrows == splits[p + 1] - splits[p] and blk.ncols == 40
        assert oracle.check_format(oracle.csc(*blk)) == 0
        parts.append(oracle.mulv(oracle.csc(*blk), x))
    np.testing.assert_array_equal(np.concatenate(parts), full)


def test_symmetric_row_block_is_transposed_column_block():
    """bench.py --gpus N builds each rank's row block as the transpose of a generated column block."""
    shape = (6, 4, 5)
    full = g.stencil3d_27pt(shape)
    r0, r1 = 40, 100
    cb = g.stencil3d_27pt(shape, col_range=(r0, r1))
    t = oracle.transpose(oracle.csc(*cb))
    rb = sharded.row_block(full, r0, r1)
    assert np.array_equal(t.colptr, rb.colptr) and np.array_equal(t.rowidx, rb.rowidx)
    assert np.array_equal(t.values, rb.values)


def test_vstack_blocks_inverts_row_block():
    """Row blocks of C concatenate back to C (what the row-partitioned SpGEMM's gather does)."""
    rng = np.random.default_rng(3)
    for m, nparts in ((random_csc(rng, 60, 45, 0.15), 4), (random_csc(rng, 7, 9, 0.5, np.complex128), 7),
                      (g.laplace3d_7pt(5), 3), (random_csc(rng, 10, 10, 0.0), 2)):
        splits = sharded.uniform_splits(m.nrows, nparts)
        blocks = [sharded.row_block(m, splits[p], splits[p + 1]) for p in range(nparts)]
        back = sharded.vstack_blocks(blocks)
        assert back.nrows == m.nrows and back.ncols == m.ncols
        assert np.array_equal(back.colptr, m.colptr) and np.array_equal(back.rowidx, m.rowidx)
        assert np.array_equal(back.values, m.values)
        assert oracle.check_format(oracle.csc(*back)) == 0


def test_row_partitioned_product_of_matrices_is_the_product():
    """mm on row blocks of A against all of B, stacked, equals mm(A, B) (oracle on both sides)."""
    rng = np.random.default_rng(8)
    a, b = random_csc(rng, 40, 30, 0.2), random_csc(rng, 30, 50, 0.2)
    want = oracle.mm(oracle.csc(*a), oracle.csc(*b))
    splits = sharded.uniform_splits(a.nrows, 3)
    blocks = []
    for p in range(3):
        c = oracle.mm(oracle.csc(*sharded.row_block(a, splits[p], splits[p + 1])), oracle.csc(*b))
        blocks.append(g.CSC(c.nrows, c.ncols, c.colptr, c.rowidx, c.values))
    got = sharded.vstack_blocks(blocks)
    assert np.array_equal(got.colptr, want.colptr) and np.array_equal(got.rowidx, want.rowidx)
    assert np.array_equal(got.values, want.values)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        gm = g.stencil3d_27pt((8, 5, 6))
        plane = 30
        splits = sharded.uniform_splits(8, world) * plane
        r0, r1 = int(splits[rank]), int(splits[rank + 1])
        blk = sharded.row_block(gm, r0, r1)
        x = g.uniform_vector(gm.ncols, seed=4)
        # handle exchange plumbing: every rank must see every rank's 64-byte token in rank order
        tokens = sharded.exchange_objects(bytes([rank]) * 64, dist)
        assert [t[0] for t in tokens] == list(range(world)) and all(len(t) == 64 for t in tokens)
        # each rank owns a slice of x; the slices are gathered (the device path reads them in place)
        mine = torch.from_numpy(np.ascontiguousarray(x[r0:r1]))
        sizes = [int(splits[p + 1] - splits[p]) for p in range(world)]
        got = [torch.empty(s, dtype=torch.float64) for s in sizes]
        dist.all_gather(got, mine) if len(set(sizes)) == 1 else dist.all_gather_object(got, mine)
        xg = np.concatenate([np.asarray(t) for t in got])
        y_local = oracle.mulv(oracle.csc(*blk), xg)
        ys = [None] * world
        dist.all_gather_object(ys, y_local)
        if rank == 0:
            want = oracle.mulv(oracle.csc(*gm), x)
            ret["ok"] = bool(np.array_equal(np.concatenate(ys), want))
    finally:
        dist.destroy_process_group()


def test_two_rank_partitioned_product_gloo():
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(2, _free_port(), ret), nprocs=2, join=True)
    assert ret.get("ok") is True
