#!/usr/bin/env python
"""bench.py -- the hot path of sparse-linear on B200, measured against its HBM roofline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg5|cfg1|cfg2|cfg3|cfg4]
                    [--impl ours|reference] [--scaling strong|weak]

Default workload (BASELINE.json configs[4], the one the metric's target is quoted on): SpMV with the
27-point stencil on a 200^3 grid (8,000,000 rows, 213,847,192 nnz, Double), x and y resident in HBM.
A step is one y <- A x.  One JSON line is printed by rank 0 (see DESIGN.md "Measurement").

  value       algorithmic bytes of all ranks / max-over-ranks device time, GB/s
  e2e         same metric through the host-pointer C-ABI call (slc_d_mulv): pinned host x -> HBM,
              product, y -> pinned host, every step
  roofline    achieved GB/s of the product kernel against MEASURED_PEAKS.json's HBM copy bandwidth
  cpu_baseline  the oracle (a single-core C port of the reference's Haskell loop) on this box's host

--impl reference times that CPU port on the same workload (the reference itself is Haskell and cannot
be built in this image: no GHC).
"""
import argparse
import ctypes as C
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "spmv_effective_hbm_gbps"
NOMINAL_HBM_GBS = 8000.0
FALLBACK_HBM_GBS = 6650.0   # /opt/skills/guides/B200_PROFILING.md, used only if MEASURED_PEAKS.json is absent


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg5", choices=["cfg1", "cfg2", "cfg3", "cfg3t", "cfg3a", "cfg4", "cfg4t", "cfg5", "cfg5mm", "cfg5t"])
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"])
    ap.add_argument("--grid", type=int, default=0, help="override the stencil grid edge (debugging)")
    ap.add_argument("--kernel", type=int, default=0, help="force SpMV kernel family (slc_set_spmv_kernel)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


# ---- workloads ---------------------------------------------------------------------------------

def build_workload(name, grid=0):
    """Returns dict(desc, matrix CSC, dtype, kind, k) -- host arrays, seeded."""
    from sparse_linear_b200 import generators as g
    if name == "cfg1":
        n = grid or 1024
        return dict(kind="spmv", m=g.laplace2d_5pt(n), dtype=np.float64, k=1,
                    desc="cfg1: 2D 5-point Laplacian %dx%d, Double SpMV (mulV)" % (n, n))
    if name == "cfg2":
        n = grid or 1024
        return dict(kind="spmm", m=g.laplace2d_5pt(n, np.complex128), dtype=np.complex128, k=64,
                    desc="cfg2: 2D 5-point Laplacian %dx%d, Complex Double SpMM k=64" % (n, n))
    if name == "cfg3":
        n = grid or 128
        return dict(kind="spgemm", m=g.laplace3d_7pt(n), dtype=np.float64, k=1,
                    desc="cfg3: SpGEMM A*A, 3D 7-point Laplacian %d^3" % n)
    if name in ("cfg3t", "cfg3a"):
        n = grid or 128
        what = "transpose" if name == "cfg3t" else "assemble"
        return dict(kind=what, m=g.laplace3d_7pt(n), dtype=np.float64, k=1,
                    desc="cfg3: %s, 3D 7-point Laplacian %d^3" % (
                        "transpose" if name == "cfg3t" else
                        "triplet assembly (every entry split in 0.25 v + 0.75 v, shuffled: 2x duplicates)", n))
    if name == "cfg5t":
        n = grid or 200
        return dict(kind="transpose", m=g.stencil3d_27pt(n), dtype=np.float64, k=1,
                    desc="cfg5: transpose, 3D 27-point stencil %d^3" % n)
    if name == "cfg4t":
        scale = grid or 24
        return dict(kind="rmat", then="transpose", scale=scale, dtype=np.float64, k=1,
                    desc="cfg4: R-MAT scale %d edge factor 16, transpose" % scale)
    if name == "cfg4":
        scale = grid or 24
        return dict(kind="rmat", scale=scale, dtype=np.float64, k=1,
                    desc="cfg4: R-MAT scale %d edge factor 16, Double SpMV" % scale)
    if name == "cfg5mm":
        n = grid or 200
        return dict(kind="spmm", m=g.stencil3d_27pt(n), dtype=np.float64, k=64,
                    desc="cfg5: 3D 27-point stencil %d^3, Double SpMM k=64" % n)
    n = grid or 200
    return dict(kind="spmv", m=g.stencil3d_27pt(n), dtype=np.float64, k=1,
                desc="cfg5: 3D 27-point stencil %d^3 (%d rows), Double SpMV (mulV)" % (n, n ** 3))


def hot_kernel(plan, k, dm):
    """Name of the kernel one step launches (what the roofline object describes)."""
    if k == 1:
        if "sell" in plan:
            return "k_spmv_sell"
        return ("k_spmv_merge" if plan.startswith("csr-merge") else "k_spmv_bal") + " + k_spmv_bal_fixup"
    if "sell" not in plan:
        return "k_spmm_rows"
    maxlen = int(plan.split("max_rowlen=")[1].split()[0])
    if maxlen > 8:
        return "k_spmm_panel"
    # rows of <= 8 entries: banded matrices take the TMA-staged window kernel unless it is switched off
    return "k_spmm_short" if os.environ.get("SLC_SPMM_WINDOW", "1") == "0" else "k_spmm_window (banded) / k_spmm_short"


def stored_bytes_note(plan, k, abytes, nnz, ms):
    """SpMV on a sliced-ELL plan with compressed indices reads fewer bytes than SURVEY.md 8(d)'s
    algorithmic count (4 B per 32 unit-stride entries instead of 4 B each): say so next to `frac`,
    which can then exceed 1, and give the rate of the bytes the format really holds."""
    if k != 1 or "unit_stride=" not in plan:
        return {}
    unit = float(plan.split("unit_stride=")[1].split("%")[0]) / 100.0
    if unit <= 0.0:
        return {}
    stored = abytes - unit * nnz * 4 + unit * nnz * 4 / 32.0
    return {"stored_bytes_per_step": int(stored), "stored_bytes_rate_gbs": stored / (ms * 1e-3) / 1e9,
            "note": "index compression: %.1f %% of the entries take their column from a 4-byte base per 32 entries, so "
                    "the kernel moves fewer bytes than the algorithmic count `achieved` is defined on (SURVEY.md 8(d)); "
                    "stored_bytes_rate_gbs is the rate of the bytes actually held in HBM" % (100.0 * unit)}


def algorithmic_bytes(kind, nrows, ncols, nnz, k, dtype):
    from sparse_linear_b200 import generators as g
    val = 16 if np.dtype(dtype).kind == "c" else 8
    if kind == "spmm":
        return g.spmm_bytes(nrows, ncols, nnz, k, val)
    return g.spmv_bytes(nrows, ncols, nnz, val)


# ---- clocks -------------------------------------------------------------------------------------

class ClockSampler:
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        names = {}
        for attr in dir(nv):
            if attr.startswith("nvmlClocksEventReason") or attr.startswith("nvmlClocksThrottleReason"):
                v = getattr(nv, attr)
                if isinstance(v, int) and v:
                    names.setdefault(v, attr.replace("nvmlClocksEventReason", "").replace("nvmlClocksThrottleReason", ""))
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, nm in names.items():
                    if r & bit and nm not in ("None", "GpuIdle", "All"):
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(0.002)

    def start(self):
        if self.nv:
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()

    def stop(self):
        self._stop.set()
        if self._thread:
            self._thread.join()
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ---- CPU baseline (oracle) ------------------------------------------------------------------------

def cpu_spmv_sample(m, dtype, budget_s, max_cols=None):
    """Times the oracle's column-scatter SpMV (the reference's axpy_ loop) on the first `cols`
    columns of the matrix.  Returns (GB/s, seconds per pass, description)."""
    import oracle
    from sparse_linear_b200 import generators as g
    ncols = m.ncols if max_cols is None else min(m.ncols, max_cols)
    nnz = int(m.colptr[ncols])
    sub = oracle.csc(m.nrows, ncols, m.colptr[:ncols + 1], m.rowidx[:nnz], m.values[:nnz])
    x = g.uniform_vector(ncols, seed=0, dtype=dtype)
    val = 16 if np.dtype(dtype).kind == "c" else 8
    nbytes = nnz * (val + 4) + (ncols + 1) * 4 + ncols * val + m.nrows * val
    oracle.mulv(sub, x)   # page in
    t0 = time.perf_counter()
    oracle.mulv(sub, x)
    one = time.perf_counter() - t0
    reps = max(1, min(50, int(budget_s / max(one, 1e-6))))
    times = []
    for _ in range(reps):
        t0 = time.perf_counter()
        oracle.mulv(sub, x)
        times.append(time.perf_counter() - t0)
    best = min(times)
    return nbytes / best / 1e9, best, "%d of %d columns (%d nnz), %d passes, best pass" % (ncols, m.ncols, nnz, reps), nbytes


def run_reference(args, rank):
    """--impl reference: the reference's CPU algorithm (oracle port; the Haskell original cannot be
    built here) on the same workload, rank 0 only."""
    if rank != 0:
        return
    wl = build_workload(args.workload if args.workload in ("cfg1", "cfg5") else "cfg5", args.grid)
    m = wl["m"]
    import oracle
    from sparse_linear_b200 import generators as g
    # bounded sample: a column block whose pass takes ~50 ms, so K+W passes finish in minutes
    frac_cols = max(1, m.ncols // 8) if m.colptr[-1] > 50_000_000 else m.ncols
    nnz = int(m.colptr[frac_cols])
    sub = oracle.csc(m.nrows, frac_cols, m.colptr[:frac_cols + 1], m.rowidx[:nnz], m.values[:nnz])
    x = g.uniform_vector(frac_cols, seed=0)
    nbytes = nnz * 12 + (frac_cols + 1) * 4 + frac_cols * 8 + m.nrows * 8
    for _ in range(args.warmup):
        oracle.mulv(sub, x)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        oracle.mulv(sub, x)
    dt = time.perf_counter() - t0
    gbs = nbytes * args.steps / dt / 1e9
    sample = "columns [0,%d) of %d (%d nnz) per step" % (frac_cols, m.ncols, nnz)
    line = {
        "impl": "reference", "metric": METRIC, "value": gbs, "unit": "GB/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f64",
        "data": "synthetic (seeded)",
        "config": {"workload": wl["desc"], "sample": sample},
        "cpu_baseline": {"value": gbs, "unit": "GB/s", "cores": 1, "kind": "port", "sample": sample},
        "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "oracle/ C port of axpy_/mulV (Data/Matrix/Sparse.hs:433-471); the reference is "
                "single-threaded Haskell and no GHC exists in this image",
    }
    print(json.dumps(line), flush=True)


# ---- ours ---------------------------------------------------------------------------------------------

def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch
    import sparse_linear_b200 as sl
    from sparse_linear_b200 import generators as g

    sl._capi.lib()   # fail loudly when the CUDA extension is missing: no fallback exists
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device")
    torch.cuda.set_device(local_rank)
    sl.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        if os.environ.get("NCCL_DEBUG", "").upper() in ("VERSION", "WARN"):
            os.environ.pop("NCCL_DEBUG")   # keeps NCCL's version banner off stdout (one JSON line)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    if world > 1 or args.gpus > 1:
        from sparse_linear_b200 import sharded
        return sharded.bench_main(args, rank, world, local_rank, dist)

    stream = torch.cuda.current_stream()
    sl.set_stream(stream.cuda_stream)
    peak, peak_src = peaks()
    wl = build_workload(args.workload, args.grid)
    dtype = wl["dtype"]
    lib = sl._capi.lib()

    if wl["kind"] == "rmat":
        n, r, c, v = g.rmat_triplets(wl["scale"], 16, seed=0)
        t0 = time.perf_counter()
        m = sl.compress(n, n, r, c, v)
        sl.synchronize()
        assemble_s = time.perf_counter() - t0
        del r, c, v
        wl["kind"] = wl.get("then", "spmv")
        wl["desc"] += " (%d nnz after duplicate merge; device assembly %.3f s incl. H2D)" % (sl.nonZero(m), assemble_s)
        host_m = None
    else:
        host_m = wl["m"]
        m = sl.Matrix(host_m.ncols, host_m.nrows, host_m.colptr, host_m.rowidx, host_m.values)
    dm = m.device()
    if args.kernel:
        dm.set_spmv_kernel(args.kernel)
    nrows, ncols, nnz = dm.dims()
    p = "z" if np.dtype(dtype).kind == "c" else "d"
    tdtype = torch.complex128 if p == "z" else torch.float64
    k = wl["k"]

    if wl["kind"] == "spgemm":
        return bench_spgemm(args, sl, m, wl, peak, peak_src)
    if wl["kind"] in ("transpose", "assemble"):
        return bench_structural(args, sl, m, host_m, wl, peak, peak_src)

    dm.prepare()
    sl.synchronize()
    # operands resident in HBM before the timed region
    if k == 1:
        xh = g.uniform_vector(ncols, seed=0, dtype=dtype)
        x = torch.from_numpy(xh).cuda()
        y = torch.empty(nrows, dtype=tdtype, device="cuda")
        fn = getattr(lib, "slc_%s_mulv_dev" % p)

        def step():
            rc = fn(dm.handle, C.c_void_p(x.data_ptr()), C.c_void_p(y.data_ptr()))
            if rc:
                raise RuntimeError(sl._capi.last_error())
    else:
        x = (torch.rand(k, ncols, dtype=torch.float64, device="cuda") * 2 - 1).to(tdtype)   # k columns, column-major
        if p == "z":
            x = torch.complex(torch.rand(k, ncols, dtype=torch.float64, device="cuda") * 2 - 1,
                              torch.rand(k, ncols, dtype=torch.float64, device="cuda") * 2 - 1)
        y = torch.empty(k, nrows, dtype=tdtype, device="cuda")
        fn = getattr(lib, "slc_%s_mulm_dev" % p)

        def step():
            rc = fn(dm.handle, C.c_void_p(x.data_ptr()), ncols, C.c_void_p(y.data_ptr()), nrows, k)
            if rc:
                raise RuntimeError(sl._capi.last_error())

    abytes = algorithmic_bytes(wl["kind"], nrows, ncols, nnz, k, dtype)
    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize()
    sampler = ClockSampler(local_rank)
    l0 = sl.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler.start()
    torch.cuda.synchronize()
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    torch.cuda.synchronize()
    clocks = sampler.stop()
    launches = sl.launch_count() - l0
    ms = e0.elapsed_time(e1) / args.steps
    value = abytes / (ms * 1e-3) / 1e9

    # end to end through the host-pointer C ABI call, pinned host buffers, every step
    e2e = None
    if not args.no_e2e:
        steps_e2e = max(3, min(args.steps, 20))
        if k == 1:
            xp = torch.from_numpy(g.uniform_vector(ncols, seed=0, dtype=dtype)).pin_memory()
            yp = torch.empty(nrows, dtype=tdtype).pin_memory()
            hfn = getattr(lib, "slc_%s_mulv" % p)

            def hstep():
                rc = hfn(dm.handle, C.c_void_p(xp.data_ptr()), C.c_void_p(yp.data_ptr()))
                if rc:
                    raise RuntimeError(sl._capi.last_error())
        else:
            xp = x.cpu().pin_memory()
            yp = torch.empty(k, nrows, dtype=tdtype).pin_memory()
            hfn = getattr(lib, "slc_%s_mulm" % p)

            def hstep():
                rc = hfn(dm.handle, C.c_void_p(xp.data_ptr()), ncols, C.c_void_p(yp.data_ptr()), nrows, k)
                if rc:
                    raise RuntimeError(sl._capi.last_error())
        for _ in range(2):
            hstep()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(steps_e2e):
            hstep()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / steps_e2e
        el = 16 if p == "z" else 8
        e2e = {"value": abytes / dt / 1e9, "unit": "GB/s", "ms_per_step": dt * 1e3, "steps": steps_e2e,
               "h2d_bytes_per_step": ncols * k * el, "d2h_bytes_per_step": nrows * k * el,
               "api": "slc_%s_%s (host pointers, pinned)" % (p, "mulv" if k == 1 else "mulm")}

    cpu = None
    if not args.no_cpu and host_m is not None:
        gbs, best, sample, _ = cpu_spmv_sample(host_m, dtype, budget_s=12.0)
        cpu = {"value": gbs, "unit": "GB/s", "cores": 1, "kind": "port", "sample": sample,
               "host_cores_available": os.cpu_count(), "ms_per_pass": best * 1e3}
        # the CPU leg is also the checker of what the timed end-to-end call produced (the oracle is used
        # nowhere else in this file's GPU arm)
        if e2e is not None and k == 1 and nnz <= 300_000_000:
            import oracle
            want = oracle.mulv(oracle.csc(*host_m), xp.numpy())
            e2e["matches_oracle_bit_exact"] = bool(np.array_equal(yp.numpy(), want))

    traffic = None
    tpath = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get(args.workload)
            traffic = traffic if isinstance(traffic, (int, float)) else None
        except Exception:
            traffic = None

    line = {
        "metric": METRIC if wl["kind"] == "spmv" else "spmm_effective_hbm_gbps",
        "value": value, "unit": "GB/s", "n_gpus": 1, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
        "dtype": "c128" if p == "z" else "f64", "data": "synthetic (seeded numpy PCG64)",
        "config": {"workload": wl["desc"], "rows": nrows, "cols": ncols, "nnz": nnz, "k": k,
                   "plan": dm.plan_info(), "algorithmic_bytes_per_step": abytes,
                   "l2": "inputs larger than L2 (%.0f MB streamed per step)" % (abytes / 1e6)
                   if abytes > 300e6 else "working set fits L2: warm-L2 number",
                   "pct_of_nominal_8TBs": 100.0 * value / NOMINAL_HBM_GBS},
        "clocks": clocks, "gpu_launches": int(launches),
        "e2e": e2e,
        "roofline": {"bound": "hbm", "achieved": value, "peak": peak, "unit": "GB/s", "frac": value / peak,
                     "traffic": traffic, "peak_source": peak_src,
                     "kernel": hot_kernel(dm.plan_info(), k, dm), **stored_bytes_note(dm.plan_info(), k, abytes, nnz, ms)},
        "cpu_baseline": cpu,
    }
    print(json.dumps(line), flush=True)


def bench_structural(args, sl, m, host_m, wl, peak, peak_src):
    """Transpose (handle -> handle) and triplet assembly (device triplets -> handle): throughput in
    stored entries per second, algorithmic bytes as SURVEY.md section 8(d)."""
    import torch
    from sparse_linear_b200 import generators as g
    steps = max(1, min(args.steps, 20))
    warm = max(args.warmup, 3) if args.warmup < 10 else 3
    nr, nc, nnz = m.device().dims()
    if wl["kind"] == "transpose":
        abytes = 2 * nnz * 12 + (nr + nc + 2) * 4
        units, unit_name = nnz, "nnz/s"
        metric = "transpose_nnz_per_s"
        kernel = "k_radix_scatter"

        def step():
            t = sl.transpose(m)
            del t
        check = None
    else:
        r, c, v = g.split_duplicates(host_m, seed=0)
        n_in = len(r)
        dr = torch.from_numpy(r.astype(np.int32)).cuda()
        dc = torch.from_numpy(c.astype(np.int32)).cuda()
        dv = torch.from_numpy(v).cuda()
        abytes = n_in * 16 + nnz * 12 + (nc + 1) * 4
        units, unit_name = n_in, "triplets/s"
        metric = "assembly_triplets_per_s"
        kernel = "k_radix_scatter"

        def step():
            t = sl.compress_dev(nr, nc, n_in, dr.data_ptr(), dc.data_ptr(), dv.data_ptr(), np.float64)
            del t
        built = sl.compress_dev(nr, nc, n_in, dr.data_ptr(), dc.data_ptr(), dv.data_ptr(), np.float64)
        check = bool(np.array_equal(built.pointers, host_m.colptr) and np.array_equal(built.indices, host_m.rowidx)
                     and np.array_equal(built.values, host_m.values))   # 0.25 v + 0.75 v == v exactly
        del built
    for _ in range(warm):
        step()
    torch.cuda.synchronize()
    sampler = ClockSampler(torch.cuda.current_device())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = sl.launch_count()
    sampler.start()
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    clocks = sampler.stop()
    ms = e0.elapsed_time(e1) / steps
    gbs = abytes / (ms * 1e-3) / 1e9
    line = {"metric": metric, "value": units / (ms * 1e-3), "unit": unit_name, "n_gpus": 1, "steps": steps,
            "warmup": warm, "ms_per_step": ms, "higher_is_better": True, "scaling": args.scaling,
            "vs_baseline": None, "dtype": "u32 keys + f64 values", "data": "synthetic (seeded numpy PCG64)",
            "config": {"workload": wl["desc"], "rows": nr, "cols": nc, "nnz": nnz, "units_per_step": units,
                       "algorithmic_bytes_per_step": abytes,
                       "l2": "inputs larger than L2" if abytes > 300e6 else "working set fits L2: warm-L2 number"},
            "clocks": clocks, "gpu_launches": int(sl.launch_count() - l0),
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak,
                         "traffic": None, "peak_source": peak_src, "kernel": kernel,
                         "note": "algorithmic bytes are the one-pass floor; an LSD radix sort moves ~20 B per "
                                 "key per 8-bit pass on top of it"}}
    if check is not None:
        line["config"]["result_equals_source_matrix_bit_exact"] = check
    print(json.dumps(line), flush=True)


def bench_spgemm(args, sl, m, wl, peak, peak_src):
    import torch
    steps = max(1, min(args.steps, 10))
    c = sl.mm(m, m)
    sl.synchronize()
    nnzc = sl.nonZero(c)
    del c
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = sl.launch_count()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(steps):
        c = sl.mm(m, m)
        del c
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    nr, nc, nnz = m.device().dims()
    floor_bytes = (2 * nnz + nnzc) * 12 + 3 * (nc + 1) * 4
    line = {"metric": "spgemm_output_nnz_per_s", "value": nnzc / (ms * 1e-3), "unit": "nnz/s", "n_gpus": 1,
            "steps": steps, "warmup": 1, "ms_per_step": ms, "higher_is_better": True, "scaling": args.scaling,
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": wl["desc"], "nnz_in": nnz, "nnz_out": nnzc},
            "gpu_launches": int(sl.launch_count() - l0),
            "roofline": {"bound": "hbm", "achieved": floor_bytes / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": floor_bytes / (ms * 1e-3) / 1e9 / peak, "traffic": None, "peak_source": peak_src}}
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
