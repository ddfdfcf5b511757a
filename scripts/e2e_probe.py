"""Where the end-to-end SpMV (slc_d_mulv, pinned host vectors) spends its step: chunk counts and the
pipeline with one leg switched off (SLC_E2E_DIAG), one process, one matrix."""
import ctypes as C
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sparse_linear_b200 as sl  # noqa: E402
from sparse_linear_b200 import generators as g  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 200
gm = g.stencil3d_27pt(n)
m = sl.Matrix(gm.ncols, gm.nrows, gm.colptr, gm.rowidx, gm.values)
dm = m.device()
dm.prepare()
lib = sl._capi.lib()
sl.set_stream(torch.cuda.current_stream().cuda_stream)
xp = torch.from_numpy(g.uniform_vector(gm.ncols, seed=0)).pin_memory()
yp = torch.empty(gm.nrows, dtype=torch.float64).pin_memory()


def step():
    rc = lib.slc_d_mulv(dm.handle, C.c_void_p(xp.data_ptr()), C.c_void_p(yp.data_ptr()))
    assert rc == 0, sl._capi.last_error()


def timed(reps=20):
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        step()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3


for diag, what in ((0, "full"), (8, "downloads read a buffer no kernel wrote"), (4, "no kernels"), (12, "no kernels, cold buffer")):
    os.environ["SLC_E2E_DIAG"] = str(diag)
    print("%s: %.3f ms" % (what, timed()), flush=True)
os.environ["SLC_E2E_DIAG"] = "0"
