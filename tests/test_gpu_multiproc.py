"""The real multi-GPU path under the oracle: one OS process per rank, CUDA IPC, the fused peer-memory
signal / wait, operand rewritten before every product (tests/mp_sharded_worker.py).  With two or more
GPUs every rank has its own device; on a one-GPU box the ranks share device 0 (IPC, flags and the
protocol are exercised all the same, the kernels time-slice)."""
import json
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _run(world, mode, steps, tmp_path, env=None):
    port = _free_port()
    procs, outs = [], []
    e = dict(os.environ)
    e.update(env or {})
    for rank in range(world):
        out = str(tmp_path / ("r%d.json" % rank))
        outs.append(out)
        procs.append(subprocess.Popen([sys.executable, os.path.join(HERE, "mp_sharded_worker.py"), str(rank), str(world),
                                       str(port), mode, str(steps), out], env=e, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT, text=True))
    logs = []
    for p in procs:
        try:
            logs.append(p.communicate(timeout=420)[0])
        except subprocess.TimeoutExpired:
            p.kill()
            logs.append("timeout")
    res = []
    for out, log in zip(outs, logs):
        assert os.path.exists(out), log[-2000:]
        res.append(json.load(open(out)))
    return res


@pytest.mark.parametrize("mode,steps,env", [("spmv", 200, {"SLC_SELL_KBASE": "2"}), ("spmv", 60, {"SLC_SELL_KBASE": "0"}),
                                            ("spmm", 40, {}), ("allgather", 40, {})])
def test_two_processes_ipc_fused_barrier_x_rewritten_every_step(sl, tmp_path, mode, steps, env):
    ndev = sl.device_count()
    world = 2
    if ndev < 2:
        steps = min(steps, 40)          # shared device: the ranks' kernels take turns
    res = _run(world, mode, steps, tmp_path, env)
    for r in res:
        assert r.get("ok"), r
        assert r["stats"][0] >= steps                                  # epochs completed on the device
    if mode != "allgather":
        assert res[0]["deps"] == 0b10 and res[1]["deps"] == 0b01       # neighbour-only dependency sets


def test_four_ranks_wait_for_neighbours_only(sl, tmp_path):
    """z-slabs over 4 ranks: rank r depends on r-1 and r+1 only (the round-1 kernel waited for all)."""
    if sl.device_count() < 4:
        pytest.skip("needs 4 GPUs")
    res = _run(4, "spmv", 100, tmp_path, {"SLC_SELL_KBASE": "2"})
    for r in res:
        assert r.get("ok"), r
    assert [r["deps"] for r in res] == [0b0010, 0b0101, 0b1010, 0b0100]


@pytest.mark.parametrize("mode", ["transpose", "assembly"])
def test_sharded_transpose_and_assembly_on_the_devices(sl, tmp_path, mode):
    """Two ranks, two GPUs, NCCL all-to-all-v of raw device arrays; blocks equal the oracle's bit for bit."""
    if sl.device_count() < 2:
        pytest.skip("NCCL needs one GPU per rank")
    res = _run(2, mode, 1, tmp_path)
    for r in res:
        assert r.get("ok"), r
