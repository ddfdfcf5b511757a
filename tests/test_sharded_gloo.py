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


def test_hstack_blocks_inverts_col_block_and_column_partitioned_product():
    """Column blocks concatenate back (hstack_blocks), and C[:, J] = A B[:, J] block by block is the
    oracle's product: the column-partitioned SpGEMM's host logic (the device test repeats it with slc_*_mm)."""
    rng = np.random.default_rng(13)
    for a, b, nparts in ((random_csc(rng, 40, 30, 0.2), random_csc(rng, 30, 50, 0.15), 4),
                         (g.laplace3d_7pt(5), g.laplace3d_7pt(5), 3),
                         (random_csc(rng, 6, 5, 0.5, np.complex128), random_csc(rng, 5, 3, 0.6, np.complex128), 5)):
        splits = sharded.uniform_splits(b.ncols, nparts)
        blocks = [sharded.col_block(b, splits[p], splits[p + 1]) for p in range(nparts)]
        back = sharded.hstack_blocks(blocks)
        assert np.array_equal(back.colptr, b.colptr) and np.array_equal(back.rowidx, b.rowidx)
        assert np.array_equal(back.values, b.values)
        want = oracle.mm(oracle.csc(*a), oracle.csc(*b))
        parts = [oracle.mm(oracle.csc(*a), oracle.csc(*blk)) for blk in blocks]
        got = sharded.hstack_blocks([g.CSC(c.nrows, c.ncols, c.colptr, c.rowidx, c.values) for c in parts])
        assert np.array_equal(got.colptr, want.colptr) and np.array_equal(got.rowidx, want.rowidx)
        assert np.array_equal(got.values, want.values)


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


def _oracle_transpose(blk):
    t = oracle.transpose(oracle.csc(*blk))
    return g.CSC(t.nrows, t.ncols, t.colptr, t.rowidx, t.values)


def test_sharded_transpose_pieces_single_process():
    """Ranks run one after the other: column blocks in, column blocks of the transpose out."""
    rng = np.random.default_rng(12)
    for m, world in ((random_csc(rng, 37, 53, 0.2), 4), (random_csc(rng, 20, 9, 0.4, np.complex128), 3),
                     (g.laplace3d_7pt(4), 2), (random_csc(rng, 8, 8, 0.0), 3)):
        want = oracle.transpose(oracle.csc(*m))
        cs, rs = sharded.uniform_splits(m.ncols, world), sharded.uniform_splits(m.nrows, world)
        sent = [sharded.transpose_pieces(_oracle_transpose(sharded.col_block(m, cs[q], cs[q + 1])), rs)
                for q in range(world)]
        for p in range(world):
            got = sharded.transpose_combine([sent[q][p] for q in range(world)])
            ref = sharded.col_block(want, rs[p], rs[p + 1])
            assert got.nrows == m.ncols and got.ncols == rs[p + 1] - rs[p]
            assert np.array_equal(got.colptr, ref.colptr) and np.array_equal(got.rowidx, ref.rowidx)
            assert np.array_equal(got.values, ref.values)


def test_sharded_compress_pieces_single_process():
    """Triplets with duplicates spread over ranks: every rank's block equals the block of the whole
    assembly; sums run in (rank, input position) order = the order of the concatenated input."""
    rng = np.random.default_rng(14)
    nr, nc, world = 23, 31, 3
    n = 400
    r, c = rng.integers(0, nr, n), rng.integers(0, nc, n)
    v = rng.integers(-5, 6, n).astype(np.float64)
    want = oracle.compress(nr, nc, r, c, v)
    cs = sharded.uniform_splits(nc, world)
    cuts = [0, 150, 150, n]          # rank 1 holds no triplets
    sent = [sharded.compress_pieces(r[cuts[q]:cuts[q + 1]], c[cuts[q]:cuts[q + 1]], v[cuts[q]:cuts[q + 1]], cs)
            for q in range(world)]
    for p in range(world):
        got = sharded.compress_combine(nr, int(cs[p + 1] - cs[p]), [sent[q][p] for q in range(world)], oracle.compress)
        ref = sharded.col_block(want, cs[p], cs[p + 1])
        assert np.array_equal(got.colptr, ref.colptr) and np.array_equal(got.rowidx, ref.rowidx)
        assert np.array_equal(got.values, ref.values)
    # an out-of-range column is reported by the owner's compress, as the reference does (Sparse.hs:192-212)
    bad = sharded.compress_pieces([0], [nc + 5], [1.0], cs)
    with pytest.raises(oracle.OracleError):
        sharded.compress_combine(nr, int(cs[-1] - cs[-2]), [bad[world - 1]], oracle.compress)


def _worker_structural(rank, world, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(30)            # same matrix on every rank
        m = random_csc(rng, 41, 29, 0.25)
        cs, rs = sharded.uniform_splits(m.ncols, world), sharded.uniform_splits(m.nrows, world)
        got = sharded.sharded_transpose(sharded.col_block(m, cs[rank], cs[rank + 1]), rs, dist,
                                        local_transpose=_oracle_transpose)
        want = sharded.col_block(oracle.transpose(oracle.csc(*m)), rs[rank], rs[rank + 1])
        ok = (np.array_equal(got.colptr, want.colptr) and np.array_equal(got.rowidx, want.rowidx)
              and np.array_equal(got.values, want.values))
        # assembly: this rank's share of the triplets (with duplicates)
        n = 300
        r, c = rng.integers(0, 41, n), rng.integers(0, 29, n)
        v = rng.integers(-5, 6, n).astype(np.float64)
        lo, hi = (0, 170) if rank == 0 else (170, n)
        blk = sharded.sharded_compress(41, r[lo:hi], c[lo:hi], v[lo:hi], cs, rank, dist, local_compress=oracle.compress)
        ref = sharded.col_block(oracle.compress(41, 29, r, c, v), cs[rank], cs[rank + 1])
        ok = ok and (np.array_equal(blk.colptr, ref.colptr) and np.array_equal(blk.rowidx, ref.rowidx)
                     and np.array_equal(blk.values, ref.values))
        oks = [None] * world
        dist.all_gather_object(oks, bool(ok))
        if rank == 0:
            ret["ok"] = all(oks)
    finally:
        dist.destroy_process_group()


def test_two_rank_transpose_and_assembly_gloo():
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker_structural, args=(2, _free_port(), ret), nprocs=2, join=True)
    assert ret.get("ok") is True


def test_sharding_helpers_edge_cases():
    """Empty blocks, more ranks than rows or columns, triplets on one rank only, a matrix without entries."""
    rng = np.random.default_rng(99)
    m = random_csc(rng, 5, 3, 0.6)
    want_t = oracle.transpose(oracle.csc(*m))
    world = 7                                   # more ranks than rows and than columns
    cs, rs = sharded.uniform_splits(m.ncols, world), sharded.uniform_splits(m.nrows, world)
    assert cs[0] == 0 and cs[-1] == m.ncols and rs[-1] == m.nrows
    sent = [sharded.transpose_pieces(_oracle_transpose(sharded.col_block(m, cs[q], cs[q + 1])), rs) for q in range(world)]
    cols = []
    for p in range(world):
        got = sharded.transpose_combine([sent[q][p] for q in range(world)])
        assert got.nrows == m.ncols and got.ncols == rs[p + 1] - rs[p]
        cols.append(got)
    # the column blocks of the transpose, side by side, are the transpose
    assert np.array_equal(np.concatenate([c.rowidx for c in cols]), want_t.rowidx)
    assert np.array_equal(np.concatenate([c.values for c in cols]), want_t.values)
    # assembly where every triplet sits on the last rank
    r, c, v = np.array([4, 0, 4, 2]), np.array([2, 0, 2, 1]), np.array([1.0, 2.0, 3.0, 4.0])
    empty = (np.zeros(0, np.int64), np.zeros(0, np.int64), np.zeros(0))
    sent = [sharded.compress_pieces(*empty, cs) for _ in range(world - 1)] + [sharded.compress_pieces(r, c, v, cs)]
    full = oracle.compress(5, 3, r, c, v)
    for p in range(world):
        got = sharded.compress_combine(5, int(cs[p + 1] - cs[p]), [sent[q][p] for q in range(world)], oracle.compress)
        ref = sharded.col_block(full, cs[p], cs[p + 1])
        assert np.array_equal(got.colptr, ref.colptr) and np.array_equal(got.rowidx, ref.rowidx)
        assert np.array_equal(got.values, ref.values)
    # a matrix with no entries at all
    z = random_csc(rng, 6, 4, 0.0)
    assert sharded.vstack_blocks([sharded.row_block(z, 0, 2), sharded.row_block(z, 2, 6)]).colptr.tolist() == [0] * 5
    assert sharded.balanced_splits(np.zeros(6), 4)[-1] == 6


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
