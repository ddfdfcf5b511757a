#!/usr/bin/env python
"""bench.py -- the hot path of sparse-linear on B200, measured against its HBM roofline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg5|cfg1|cfg2|cfg3|cfg3t|cfg3a|cfg4|cfg4t|cfg5mm|cfg5t]
                    [--impl ours|reference] [--scaling strong|weak] [--suite auto|none]

Default workload (BASELINE.json configs[4], the one the metric's target is quoted on): SpMV with the
27-point stencil on a 200^3 grid (8,000,000 rows, 213,847,192 nnz, Double), x and y resident in HBM.
A step is one y <- A x.  One JSON line is printed by rank 0 (see DESIGN.md "Measurement"); with the
default workload the line also carries `suite`: one entry per other op / config of BASELINE.json
(cfg 1 cold-L2 / warm / CUDA-graph, cfg 2 SpMM and the through-FEAST call, cfg 3 SpGEMM / transpose /
assembly, cfg 4 at scale 24, cfg 5 SpMM and transpose), each with its own value, clocks, roofline, e2e,
cpu_baseline and an oracle verdict on what the timed call produced.

  value       algorithmic bytes of all ranks / max-over-ranks device time, GB/s
  e2e         same metric through the host-pointer C-ABI call: pinned host operands -> HBM, op,
              result -> pinned host, every step
  roofline    achieved GB/s of the dominant kernel against MEASURED_PEAKS.json's HBM copy bandwidth
  cpu_baseline  the oracle (a single-core C port of the reference's Haskell loops) on this box's host

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
import traceback

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "spmv_effective_hbm_gbps"
NOMINAL_HBM_GBS = 8000.0
FALLBACK_HBM_GBS = 6650.0   # /opt/skills/guides/B200_PROFILING.md, used only if MEASURED_PEAKS.json is absent
WORKLOADS = ["cfg1", "cfg2", "cfg3", "cfg3t", "cfg3a", "cfg4", "cfg4t", "cfg5", "cfg5mm", "cfg5t"]


def parse_args(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg5", choices=WORKLOADS)
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"])
    ap.add_argument("--suite", default="auto", choices=["auto", "none"],
                    help="auto: the default workload also runs the other configs and reports them under `suite`")
    ap.add_argument("--grid", type=int, default=0, help="override the stencil grid edge / R-MAT scale (debugging)")
    ap.add_argument("--kernel", type=int, default=0, help="force SpMV kernel family (slc_set_spmv_kernel)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline legs")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="multi-GPU: launch every product from the host instead of one CUDA graph")
    ap.add_argument("--thorough", action="store_true", help="slower CPU legs (literal mm at 48^3)")
    return ap.parse_args(argv)


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


def traffic_of(key):
    """Per-launch DRAM bytes of the dominant kernel from the committed ncu --set full captures."""
    tpath = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    try:
        v = json.load(open(tpath)).get(key)
        return v if isinstance(v, (int, float)) else None
    except Exception:
        return None


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

    def _sample(self, names):
        nv = self.nv
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

    def _loop(self):
        nv = self.nv
        names = {}
        for attr in dir(nv):
            if attr.startswith("nvmlClocksEventReason") or attr.startswith("nvmlClocksThrottleReason"):
                v = getattr(nv, attr)
                if isinstance(v, int) and v:
                    names.setdefault(v, attr.replace("nvmlClocksEventReason", "").replace("nvmlClocksThrottleReason", ""))
        # NVML queries take driver locks that the timed thread's allocations and synchronisations also
        # need: at one sample per 2 ms the assembly step (two host synchronisations per call) measured
        # anywhere between 2.1 and 50 ms; at 25 ms the perturbation is gone and a 50 ms region still has
        # samples at its start, inside and at its end
        period = float(os.environ.get("SLC_BENCH_CLOCK_PERIOD", "0.025"))
        while not self._stop.is_set():
            self._sample(names)
            self._stop.wait(period)
        self._sample(names)

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


class Timer:
    """CUDA-event timing of K steps on torch's current stream (the stream the library launches on),
    clocks sampled during the region, kernels counted by the library."""

    def __init__(self, sl, device_index=0):
        import torch
        self.torch, self.sl, self.dev = torch, sl, device_index
        self.stream = torch.cuda.current_stream()
        self._flush = None

    def flush_l2(self):
        """Write a buffer twice the size of L2 so the next step starts with a cold cache."""
        if self._flush is None:
            self._flush = self.torch.empty(256 << 20, dtype=self.torch.uint8, device="cuda")
        self._flush.add_(1)

    def run(self, step, steps, warmup, cold=False, min_seconds=0.0):
        """step(i) enqueues step i.  Returns dict(ms, clocks, launches).  cold: L2 flushed before every
        step, each step bracketed by its own events (the flush is not timed).  min_seconds: for very short
        steps the region is repeated until the clock sampler has something to see; ms stays per step."""
        torch = self.torch
        for i in range(max(warmup, 3)):
            step(i)
        torch.cuda.synchronize()
        sampler = ClockSampler(self.dev)
        l0 = self.sl.launch_count()
        sampler.start()
        torch.cuda.synchronize()
        t_host = time.perf_counter()
        total_ms, done, per_repeat = 0.0, 0, []
        while True:
            before = total_ms
            if cold:
                evs = []
                for i in range(steps):
                    self.flush_l2()
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record(self.stream)
                    step(i)
                    b.record(self.stream)
                    evs.append((a, b))
                torch.cuda.synchronize()
                total_ms += sum(a.elapsed_time(b) for a, b in evs)
            else:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(self.stream)
                for i in range(steps):
                    step(i)
                e1.record(self.stream)
                torch.cuda.synchronize()
                total_ms += e0.elapsed_time(e1)
            done += steps
            per_repeat.append((total_ms - before) / steps)
            if time.perf_counter() - t_host >= min_seconds and (min_seconds == 0.0 or len(per_repeat) >= 3):
                break
        clocks = sampler.stop()
        launches = self.sl.launch_count() - l0
        # steps with host synchronisations inside (mm, transpose, compress) are exposed to whatever else takes
        # a driver lock on the box (the clock sampler's own NVML queries among them): such regions are repeated
        # at least three times and the MEDIAN repeat is reported, with the spread beside it
        ms = statistics.median(per_repeat) if len(per_repeat) >= 3 else total_ms / done
        return {"ms": ms, "clocks": clocks, "launches": int(round(launches * steps / done)), "steps": steps,
                "repeats_for_clock_sampling": done // steps,
                "ms_per_repeat": ([round(x, 5) for x in per_repeat] if len(per_repeat) <= 12 else
                                  {"n": len(per_repeat), "min": round(min(per_repeat), 5), "median": round(ms, 5),
                                   "max": round(max(per_repeat), 5)})}


def roofline(achieved_gbs, kernel, traffic_key=None, **extra):
    peak, src = peaks()
    r = {"bound": "hbm", "achieved": achieved_gbs, "peak": peak, "unit": "GB/s", "frac": achieved_gbs / peak,
         "traffic": traffic_of(traffic_key) if traffic_key else None, "peak_source": src, "kernel": kernel}
    r.update(extra)
    return r


def lib_call(sl, rc):
    if rc:
        raise RuntimeError(sl._capi.last_error())


# ---- SpMV ------------------------------------------------------------------------------------------

def stored_bytes_note(plan, abytes, nnz, ms):
    """SpMV on a sliced-ELL plan with compressed indices reads fewer bytes than SURVEY.md 8(d)'s
    algorithmic count (4 B per 32 unit-stride entries instead of 4 B each): say so next to `frac`,
    which can then exceed 1, and give the rate of the bytes the format really holds."""
    if "unit_stride=" not in plan:
        return {}
    unit = float(plan.split("unit_stride=")[1].split("%")[0]) / 100.0
    if unit <= 0.0:
        return {}
    stored = abytes - unit * nnz * 4 + unit * nnz * 4 / 32.0
    peak, _ = peaks()
    return {"stored_bytes_per_step": int(stored), "stored_bytes_rate_gbs": stored / (ms * 1e-3) / 1e9,
            "stored_bytes_frac_of_peak": stored / (ms * 1e-3) / 1e9 / peak,
            "note": "index compression: %.1f %% of the entries take their column from a 4-byte base per 32 entries, so "
                    "the kernel moves fewer bytes than the algorithmic count `achieved` is defined on (SURVEY.md 8(d)); "
                    "stored_bytes_frac_of_peak is the hardware fraction, `frac` the 8(d) figure" % (100.0 * unit)}


def spmv_kernel_name(plan):
    if plan.startswith("sell32"):
        return "k_spmv_sell"
    return ("k_spmv_merge" if plan.startswith("csr-merge") else "k_spmv_bal") + " + k_spmv_bal_fixup"


def cpu_spmv(om, x, budget_s):
    """The oracle's column-scatter SpMV (the reference's axpy_ loop) on one host core: best of a few
    passes over the whole matrix.  Returns (seconds per pass, passes, result of the last pass)."""
    import oracle
    t0 = time.perf_counter()
    y = oracle.mulv(om, x)
    one = time.perf_counter() - t0
    reps = max(1, min(20, int(budget_s / max(one, 1e-6))))
    best = one
    for _ in range(reps):
        t0 = time.perf_counter()
        y = oracle.mulv(om, x)
        best = min(best, time.perf_counter() - t0)
    return best, reps + 1, y


def entry_spmv(args, sl, tm, name, desc, host_m=None, dev_m=None, dtype=np.float64, graph_batch=False, cold=False,
               traffic_key=None, cpu_budget=6.0):
    """y <- A x with x, y resident; x rotates through 3 buffers so no step can lean on the previous
    step's operand sitting in L2.  The last timed product is downloaded and compared with the oracle."""
    import torch
    import oracle
    from sparse_linear_b200 import generators as g
    lib = sl._capi.lib()
    m = dev_m if dev_m is not None else sl.Matrix(host_m.ncols, host_m.nrows, host_m.colptr, host_m.rowidx, host_m.values)
    dm = m.device()
    if args.kernel:
        dm.set_spmv_kernel(args.kernel)
    dm.prepare()
    sl.synchronize()
    nrows, ncols, nnz = dm.dims()
    p = "z" if np.dtype(dtype).kind == "c" else "d"
    tdt = torch.complex128 if p == "z" else torch.float64
    val = 16 if p == "z" else 8
    xh = [g.uniform_vector(ncols, seed=s, dtype=dtype) for s in range(3)]
    xs = [torch.from_numpy(a).cuda() for a in xh]
    y = torch.empty(nrows, dtype=tdt, device="cuda")
    fn = getattr(lib, "slc_%s_mulv_dev" % p)

    def step(i):
        lib_call(sl, fn(dm.handle, C.c_void_p(xs[i % 3].data_ptr()), C.c_void_p(y.data_ptr())))
    abytes = g.spmv_bytes(nrows, ncols, nnz, val)
    steps = args.steps
    r = tm.run(step, steps, args.warmup, min_seconds=0.05)
    ms = r["ms"]
    plan = dm.plan_info()
    value = abytes / (ms * 1e-3) / 1e9
    entry = {"name": name, "metric": METRIC, "value": value, "unit": "GB/s", "ms_per_step": ms, "steps": steps,
             "dtype": "c128" if p == "z" else "f64", "clocks": r["clocks"], "gpu_launches": r["launches"], "ms_per_repeat": r.get("ms_per_repeat"),
             "config": {"workload": desc, "rows": nrows, "cols": ncols, "nnz": nnz, "plan": plan,
                        "algorithmic_bytes_per_step": abytes, "operand": "x rotates through 3 device buffers",
                        "l2": "inputs larger than L2 (%.0f MB streamed per step)" % (abytes / 1e6)
                        if abytes > 300e6 else "working set fits L2: warm-L2 number",
                        "pct_of_nominal_8TBs": 100.0 * value / NOMINAL_HBM_GBS},
             "roofline": roofline(value, spmv_kernel_name(plan), traffic_key, **stored_bytes_note(plan, abytes, nnz, ms))}
    last = (steps - 1) % 3
    y_dev = y.cpu().numpy()
    if cold:
        rc = tm.run(step, min(steps, 50), 3, cold=True)
        entry["cold_l2"] = {"ms_per_step": rc["ms"], "value": abytes / (rc["ms"] * 1e-3) / 1e9, "unit": "GB/s",
                            "frac": abytes / (rc["ms"] * 1e-3) / 1e9 / peaks()[0], "clocks": rc["clocks"],
                            "how": "256 MB written before every step; each step timed by its own event pair"}
    if graph_batch:
        # launch overhead made visible: the same K products captured once in a CUDA graph and replayed
        cap = torch.cuda.Stream()
        sl.set_stream(cap.cuda_stream)
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.stream(cap):
            step(0)
            cap.synchronize()
            with torch.cuda.graph(gr, stream=cap):
                for i in range(20):
                    step(i)
        sl.set_stream(tm.stream.cuda_stream)
        rg = tm.run(lambda i: gr.replay(), max(5, steps // 20), 3, min_seconds=0.05)
        entry["cuda_graph"] = {"products_per_graph": 20, "ms_per_product": rg["ms"] / 20,
                               "value": abytes / (rg["ms"] / 20 * 1e-3) / 1e9, "unit": "GB/s", "clocks": rg["clocks"],
                               "frac": abytes / (rg["ms"] / 20 * 1e-3) / 1e9 / peaks()[0]}
    # end to end through the host-pointer C ABI call, pinned host buffers, every step
    xp = yp = None
    if not args.no_e2e:
        steps_e2e = max(3, min(steps, 20))
        xp = torch.from_numpy(xh[0]).pin_memory()
        yp = torch.empty(nrows, dtype=tdt).pin_memory()
        hfn = getattr(lib, "slc_%s_mulv" % p)

        def hstep():
            lib_call(sl, hfn(dm.handle, C.c_void_p(xp.data_ptr()), C.c_void_p(yp.data_ptr())))
        for _ in range(2):
            hstep()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(steps_e2e):
            hstep()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / steps_e2e
        entry["e2e"] = {"value": abytes / dt / 1e9, "unit": "GB/s", "ms_per_step": dt * 1e3, "steps": steps_e2e,
                        "h2d_bytes_per_step": ncols * val, "d2h_bytes_per_step": nrows * val,
                        "api": "slc_%s_mulv (host pointers, pinned)" % p}
        # the same call on the memory an unchanged caller of the reference owns: plain (pageable) arrays
        xg = np.array(xh[0], copy=True)
        yg = np.zeros(nrows, dtype=xg.dtype)

        def gstep():
            lib_call(sl, hfn(dm.handle, xg.ctypes.data_as(C.c_void_p), yg.ctypes.data_as(C.c_void_p)))
        dtg, repsg = timed_host(gstep, 5)
        entry["e2e_pageable"] = {"value": abytes / dtg / 1e9, "unit": "GB/s", "ms_per_step": dtg * 1e3, "steps": len(repsg),
                                 "ms_per_repeat": repsg, "api": "slc_%s_mulv (host pointers, pageable numpy arrays, reused)" % p,
                                 "same_result_as_pinned": bool(np.array_equal(yg, yp.numpy()))}
    # CPU leg: the baseline, and the checker of what the timed calls produced
    if not args.no_cpu:
        if host_m is None:
            host_m = g.CSC(m.nrows, m.ncols, m.pointers, m.indices, m.values)     # device-assembled: bring it back
        om = oracle.csc(*host_m)
        best, passes, want = cpu_spmv(om, xh[0], cpu_budget)
        entry["cpu_baseline"] = {"value": abytes / best / 1e9, "unit": "GB/s", "cores": 1, "kind": "port",
                                 "sample": "whole matrix, %d passes, best pass" % passes,
                                 "host_cores_available": os.cpu_count(), "ms_per_pass": best * 1e3}
        exact = plan.startswith("sell32")     # one thread per row: the reference's summation order
        want_last = want if last == 0 else oracle.mulv(om, xh[last])
        entry["parity"] = {"oracle": "oracle.mulv (Sparse.hs:433-471)", "call": "slc_%s_mulv_dev, last timed step" % p,
                           "bit_exact": bool(np.array_equal(y_dev, want_last)),
                           "rel_err": float(np.linalg.norm(y_dev - want_last) / max(np.linalg.norm(want_last), 1e-300)),
                           "required": "bit-exact" if exact else "1e-12"}
        entry["parity"]["ok"] = entry["parity"]["bit_exact"] if exact else entry["parity"]["rel_err"] <= 1e-12
        if yp is not None:
            got = yp.numpy()
            entry["e2e"]["matches_oracle_bit_exact"] = bool(np.array_equal(got, want))
            entry["e2e"]["rel_err"] = float(np.linalg.norm(got - want) / max(np.linalg.norm(want), 1e-300))
    return entry, m


# ---- SpMM -----------------------------------------------------------------------------------------

def spmm_kernel_name(plan):
    if not plan.startswith("sell32"):
        return "k_spmm_rows"
    maxlen = int(plan.split("max_rowlen=")[1].split()[0])
    if maxlen > 8:
        return "k_spmm_planes" if "brick=1" in plan else "k_spmm_panel"
    return "k_spmm_short" if os.environ.get("SLC_SPMM_WINDOW", "1") == "0" else "k_spmm_window (TMA)"


def entry_spmm(args, sl, tm, name, desc, host_m, dtype, k, check_cols, traffic_key=None, m=None, feast_leg=False):
    import torch
    import oracle
    from sparse_linear_b200 import generators as g
    lib = sl._capi.lib()
    if m is None:
        m = sl.Matrix(host_m.ncols, host_m.nrows, host_m.colptr, host_m.rowidx, host_m.values)
    dm = m.device()
    dm.prepare()
    nrows, ncols, nnz = dm.dims()
    p = "z" if np.dtype(dtype).kind == "c" else "d"
    tdt = torch.complex128 if p == "z" else torch.float64
    val = 16 if p == "z" else 8
    gen = torch.Generator(device="cuda")
    gen.manual_seed(1234)

    def rnd():
        a = torch.rand(k, ncols, dtype=torch.float64, device="cuda", generator=gen) * 2 - 1
        if p == "z":
            a = torch.complex(a, torch.rand(k, ncols, dtype=torch.float64, device="cuda", generator=gen) * 2 - 1)
        return a
    x = rnd()                                   # k columns, column-major
    y = torch.empty(k, nrows, dtype=tdt, device="cuda")
    fn = getattr(lib, "slc_%s_mulm_dev" % p)

    def step(i):
        lib_call(sl, fn(dm.handle, C.c_void_p(x.data_ptr()), ncols, C.c_void_p(y.data_ptr()), nrows, k, None))
    abytes = g.spmm_bytes(nrows, ncols, nnz, k, val)
    steps = max(3, min(args.steps, 20))
    r = tm.run(step, steps, 3, min_seconds=0.05)
    ms = r["ms"]
    value = abytes / (ms * 1e-3) / 1e9
    plan = dm.plan_info()
    entry = {"name": name, "metric": "spmm_effective_hbm_gbps", "value": value, "unit": "GB/s", "ms_per_step": ms,
             "steps": steps, "dtype": "c128" if p == "z" else "f64", "clocks": r["clocks"], "gpu_launches": r["launches"], "ms_per_repeat": r.get("ms_per_repeat"),
             "config": {"workload": desc, "rows": nrows, "cols": ncols, "nnz": nnz, "k": k, "plan": plan,
                        "algorithmic_bytes_per_step": abytes, "l2": "inputs larger than L2 (%.0f MB per step)" % (abytes / 1e6),
                        "pct_of_nominal_8TBs": 100.0 * value / NOMINAL_HBM_GBS},
             "roofline": roofline(value, spmm_kernel_name(plan), traffic_key)}
    # parity of the timed device call: chosen columns against the reference loop
    om = oracle.csc(*host_m)
    cols = sorted(set(check_cols))
    xcols = np.asfortranarray(x[cols].cpu().numpy().T)
    ycols = y[cols].cpu().numpy().T
    want = t_cpu = None
    if not args.no_cpu:
        t0 = time.perf_counter()
        want = oracle.mulm(om, xcols)
        t_cpu = (time.perf_counter() - t0) / len(cols)
        entry["parity"] = {"oracle": "oracle.mulm (Feast.hs:203-208: set 0, axpy_ per column)", "call": "slc_%s_mulm_dev" % p,
                           "columns_checked": cols, "bit_exact": bool(np.array_equal(ycols, want)), "required": "bit-exact"}
        entry["parity"]["ok"] = entry["parity"]["bit_exact"]
        entry["cpu_baseline"] = {"value": abytes / (t_cpu * k) / 1e9, "unit": "GB/s", "cores": 1, "kind": "port",
                                 "sample": "%d of %d columns timed (%.1f ms each); the reference serves the columns one by "
                                           "one, so k columns cost k times one" % (len(cols), k, t_cpu * 1e3),
                                 "host_cores_available": os.cpu_count()}
    if not args.no_e2e:
        xp = x.cpu().pin_memory()
        yp = torch.empty(k, nrows, dtype=tdt).pin_memory()
        hfn = getattr(lib, "slc_%s_mulm" % p)

        def hstep():
            lib_call(sl, hfn(dm.handle, C.c_void_p(xp.data_ptr()), ncols, C.c_void_p(yp.data_ptr()), nrows, k, None))
        hstep()
        torch.cuda.synchronize()
        n_e2e = 3
        t0 = time.perf_counter()
        for _ in range(n_e2e):
            hstep()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / n_e2e
        entry["e2e"] = {"value": abytes / dt / 1e9, "unit": "GB/s", "ms_per_step": dt * 1e3, "steps": n_e2e,
                        "h2d_bytes_per_step": ncols * k * val, "d2h_bytes_per_step": nrows * k * val,
                        "api": "slc_%s_mulm (host pointers, pinned; pipelined by column groups)" % p}
        if want is not None:
            entry["e2e"]["matches_oracle_bit_exact"] = bool(np.array_equal(yp[cols].numpy().T, want))
        # the same call on plain (pageable) blocks, reused from call to call as FEAST reuses its work arrays
        xg = np.array(xp.numpy(), copy=True)
        yg = np.zeros_like(yp.numpy())

        def gstep():
            lib_call(sl, hfn(dm.handle, xg.ctypes.data_as(C.c_void_p), ncols, yg.ctypes.data_as(C.c_void_p), nrows, k, None))
        dtg, repsg = timed_host(gstep, 3)
        entry["e2e_pageable"] = {"value": abytes / dtg / 1e9, "unit": "GB/s", "ms_per_step": dtg * 1e3, "steps": len(repsg),
                                 "ms_per_repeat": repsg, "api": "slc_%s_mulm (host pointers, pageable numpy blocks, reused)" % p,
                                 "same_result_as_pinned": bool(np.array_equal(yg, yp.numpy()))}
        del xg, yg
        if feast_leg:
            entry["through_feast"] = feast_leg_run(sl, m, host_m, dtype, k, xp, want, cols, abytes)
        del xp, yp
    return entry


def feast_leg_run(sl, m, host_m, dtype, m0, xp, want, cols, abytes):
    """cfg 2 as FEAST drives it (Feast.hs:189-208): eigenvectors / work1 are host arrays of m0 columns,
    ijob 30 with fpm[23], fpm[24] picks the active range, one slc_z_mulm per call.  Reported beside the
    device-resident number, never mixed with it."""
    from sparse_linear_b200 import feast
    ws = feast.FeastWorkspace(host_m.ncols, m0, dtype)
    ws.eigenvectors[...] = xp.numpy().T
    out = {"api": "FeastWorkspace.multiplyWork -> slc_%s_mulm on FEAST's own host blocks (pinned)" % ws.p, "calls": []}
    for first, count in ((1, m0), (1 + m0 // 4, m0 // 2)):
        ws.fpm[23], ws.fpm[24] = first, count
        ws.multiplyWork(m)
        t0 = time.perf_counter()
        for _ in range(2):
            ws.multiplyWork(m)
        dt = (time.perf_counter() - t0) / 2
        rec = {"fpm23": first, "fpm24": count, "ms_per_call": dt * 1e3, "h2d_bytes": count * host_m.ncols * ws.dtype.itemsize,
               "d2h_bytes": count * host_m.nrows * ws.dtype.itemsize}
        if count == m0:
            rec["value"] = abytes / dt / 1e9
            rec["unit"] = "GB/s"
            if want is not None:
                rec["matches_oracle_bit_exact"] = bool(np.array_equal(ws.work1[:, cols], want))
        out["calls"].append(rec)
    ws.close()
    return out


# ---- SpGEMM, transpose, assembly ------------------------------------------------------------------------

def same_matrix(mat, o):
    return bool(mat.nrows == o.nrows and mat.ncols == o.ncols and np.array_equal(mat.pointers, o.colptr)
                and np.array_equal(mat.indices, o.rowidx) and np.array_equal(mat.values, o.values))


def timed_host(hstep, repeats=5):
    """End-to-end timing of a handle-producing call on host arrays: one untimed call, then `repeats` timed
    ones; the median is the figure (round 2 published a single un-repeated step)."""
    hstep()
    ts = []
    for _ in range(repeats):
        t0 = time.perf_counter()
        hstep()
        ts.append(time.perf_counter() - t0)
    return float(np.median(ts)), [round(t * 1e3, 3) for t in ts]


def entry_spgemm(args, sl, tm, host_m, n):
    import oracle
    from sparse_linear_b200 import generators as g
    m = sl.Matrix(host_m.ncols, host_m.nrows, host_m.colptr, host_m.rowidx, host_m.values)
    m.device()
    keep = {}

    def step(i):
        keep["c"] = sl.mm(m, m)
    steps = max(3, min(args.steps, 20))
    r = tm.run(step, steps, 3, min_seconds=0.05)
    ms = r["ms"]
    c = keep["c"]
    nnzc = sl.nonZero(c)
    nr, nc, nnz = m.device().dims()
    floor_bytes = (2 * nnz + nnzc) * 12 + 3 * (nc + 1) * 4
    gbs = floor_bytes / (ms * 1e-3) / 1e9
    entry = {"name": "cfg3_spgemm", "metric": "spgemm_output_nnz_per_s", "value": nnzc / (ms * 1e-3), "unit": "nnz/s",
             "ms_per_step": ms, "steps": steps, "dtype": "f64", "clocks": r["clocks"], "gpu_launches": r["launches"], "ms_per_repeat": r.get("ms_per_repeat"),
             "config": {"workload": "cfg3: SpGEMM A*A, 3D 7-point Laplacian %d^3 (symbolic + numeric, handles resident)" % n,
                        "nnz_in": nnz, "nnz_out": nnzc, "byte_floor": floor_bytes},
             "roofline": roofline(gbs, "k_mm_merge (numeric)", "cfg3",
                                  note="achieved = one-pass byte floor (nnzA+nnzB+nnzC)*(val+idx)+3*(m+1)*ptr over the step")}
    if not args.no_e2e:
        def hstep():
            a = sl.Matrix(host_m.ncols, host_m.nrows, host_m.colptr, host_m.rowidx, host_m.values)
            cc = sl.mm(a, a)
            return cc.pointers, cc.indices, cc.values            # download
        dt, reps = timed_host(hstep)
        entry["e2e"] = {"value": nnzc / dt, "unit": "nnz/s", "ms_per_step": dt * 1e3, "steps": len(reps), "ms_per_repeat": reps,
                        "h2d_bytes_per_step": int(nnz * 12 + (nc + 1) * 4), "d2h_bytes_per_step": int(nnzc * 12 + (nc + 1) * 4),
                        "api": "slc_d_upload + slc_d_mm + slc_d_download (pageable host arrays, as the shim passes them; "
                               "results into fresh numpy buffers); median of the repeats"}
    if not args.no_cpu:
        om = oracle.csc(*host_m)
        t0 = time.perf_counter()
        want = oracle.mm(om, om)
        t_fast = time.perf_counter() - t0
        entry["parity"] = {"oracle": "oracle.mm (Sparse.hs:691-702, ScatterGather.hs:37-147; touched-list accumulator)",
                           "structure_bit_exact": bool(np.array_equal(c.pointers, want.colptr) and np.array_equal(c.indices, want.rowidx)),
                           "values_bit_exact": bool(np.array_equal(c.values, want.values)), "required": "structure and values bit-exact"}
        entry["parity"]["ok"] = entry["parity"]["structure_bit_exact"] and entry["parity"]["values_bit_exact"]
        lit = []
        for q in ((24, 32, 48) if args.thorough else (24, 32)):
            a = g.laplace3d_7pt(q)
            oa = oracle.csc(*a)
            t0 = time.perf_counter()
            cl = oracle.mm(oa, oa, literal=True)
            lit.append({"grid": "%d^3" % q, "seconds": time.perf_counter() - t0, "nnz_out": int(cl.colptr[-1])})
        big = lit[-1]
        q = int(big["grid"].split("^")[0])
        extrap = big["seconds"] * (n / q) ** 6      # reset + 3 scans of length nrows per output column: O(n^6) in the grid edge
        entry["cpu_baseline"] = {"value": nnzc / t_fast, "unit": "nnz/s", "cores": 1, "kind": "port",
                                 "sample": "output-identical fast restatement (touched-list reset) at %d^3, one pass: %.2f s" % (n, t_fast),
                                 "literal_restatement": lit,
                                 "literal_extrapolated_to_%d^3" % n: {"seconds": extrap, "value": nnzc / extrap, "unit": "nnz/s",
                                                                      "EXTRAPOLATED": "quadratic law t ~ rows^2 from the largest timed grid; never measured"},
                                 "host_cores_available": os.cpu_count()}
    return entry


def entry_transpose(args, sl, tm, name, desc, m, host_m, traffic_key=None, oracle_check=True, symmetric=False):
    import oracle
    keep = {}

    def step(i):
        keep["t"] = sl.transpose(m)
    steps = max(3, min(args.steps, 20))
    r = tm.run(step, steps, 3, min_seconds=0.05)
    ms = r["ms"]
    nr, nc, nnz = m.device().dims()
    abytes = 2 * nnz * 12 + (nr + nc + 2) * 4
    gbs = abytes / (ms * 1e-3) / 1e9
    entry = {"name": name, "metric": "transpose_nnz_per_s", "value": nnz / (ms * 1e-3), "unit": "nnz/s", "ms_per_step": ms,
             "steps": steps, "dtype": "u32 keys + f64 values", "clocks": r["clocks"], "gpu_launches": r["launches"], "ms_per_repeat": r.get("ms_per_repeat"),
             "config": {"workload": desc, "rows": nr, "cols": nc, "nnz": nnz, "algorithmic_bytes_per_step": abytes},
             "roofline": roofline(gbs, "k_radix_scatter", traffic_key,
                                  note="algorithmic bytes are the two-pass floor 2*nnz*(val+idx); an LSD radix sort moves more")}
    t = keep["t"]
    if host_m is None and not (args.no_e2e and args.no_cpu):
        from sparse_linear_b200 import generators as g
        host_m = g.CSC(nr, nc, m.pointers, m.indices, m.values)       # the handle's arrays, downloaded once
    if not args.no_e2e and host_m is not None:
        def hstep():
            a = sl.Matrix(host_m.ncols, host_m.nrows, host_m.colptr, host_m.rowidx, host_m.values)
            tt = sl.transpose(a)
            return tt.pointers, tt.indices, tt.values
        dt, reps = timed_host(hstep, 5 if nnz < 100_000_000 else 3)
        entry["e2e"] = {"value": nnz / dt, "unit": "nnz/s", "ms_per_step": dt * 1e3, "steps": len(reps), "ms_per_repeat": reps,
                        "h2d_bytes_per_step": int(nnz * 12 + (nc + 1) * 4), "d2h_bytes_per_step": int(nnz * 12 + (nr + 1) * 4),
                        "api": "slc_d_upload + slc_d_transpose + slc_d_download (pageable host arrays; results into fresh "
                               "numpy buffers); median of the repeats"}
    if not args.no_cpu:
        par = {"required": "structure and values bit-exact"}
        if oracle_check and host_m is not None:
            om = oracle.csc(*host_m)
            t0 = time.perf_counter()
            want = oracle.transpose(om)
            dt = time.perf_counter() - t0
            par.update(oracle="oracle.transpose (Sparse.hs:301-329)", bit_exact=same_matrix(t, want))
            par["ok"] = par["bit_exact"]
            if symmetric:
                par["fixed_point_bit_exact"] = bool(t == m)
                par["ok"] = par["ok"] and par["fixed_point_bit_exact"]
            entry["cpu_baseline"] = {"value": nnz / dt, "unit": "nnz/s", "cores": 1, "kind": "port",
                                     "sample": "whole matrix, one pass: %.2f s" % dt, "host_cores_available": os.cpu_count()}
        else:
            # full-size property instead of a 20 s oracle pass: transposing twice gives the matrix back bit for bit
            # (and a symmetric matrix is its own transpose)
            back = sl.transpose(t)
            par.update(oracle="property: transpose (transpose A) == A" + (" and transpose A == A (symmetric)" if symmetric else ""),
                       round_trip_bit_exact=bool(back == m))
            par["ok"] = par["round_trip_bit_exact"]
            if symmetric:
                par["fixed_point_bit_exact"] = bool(t == m)
                par["ok"] = par["ok"] and par["fixed_point_bit_exact"]
        entry["parity"] = par
    return entry


def entry_assembly(args, sl, tm, host_m, n):
    import torch
    import oracle
    from sparse_linear_b200 import generators as g
    r_, c_, v_ = g.split_duplicates(host_m, seed=0)
    n_in = len(r_)
    r32, c32 = r_.astype(np.int32), c_.astype(np.int32)
    dr, dc, dv = torch.from_numpy(r32).cuda(), torch.from_numpy(c32).cuda(), torch.from_numpy(v_).cuda()
    nr, nc, nnz = host_m.nrows, host_m.ncols, int(host_m.colptr[-1])
    keep = {}

    def step(i):
        keep["a"] = sl.compress_dev(nr, nc, n_in, dr.data_ptr(), dc.data_ptr(), dv.data_ptr(), np.float64)
    steps = max(3, min(args.steps, 20))
    r = tm.run(step, steps, 3, min_seconds=0.05)
    ms = r["ms"]
    abytes = n_in * 16 + nnz * 12 + (nc + 1) * 4
    gbs = abytes / (ms * 1e-3) / 1e9
    built = keep["a"]
    entry = {"name": "cfg3_assembly", "metric": "assembly_triplets_per_s", "value": n_in / (ms * 1e-3), "unit": "triplets/s",
             "ms_per_step": ms, "steps": steps, "dtype": "u64 keys + f64 values", "clocks": r["clocks"], "gpu_launches": r["launches"], "ms_per_repeat": r.get("ms_per_repeat"),
             "config": {"workload": "cfg3: triplet assembly, 3D 7-point Laplacian %d^3, every entry split in 0.25 v + 0.75 v and "
                                    "shuffled (2x duplicates); triplets resident in HBM (slc_d_compress_dev)" % n,
                        "triplets_in": n_in, "nnz_out": nnz, "algorithmic_bytes_per_step": abytes},
             "roofline": roofline(gbs, "k_coarse_scatter", "cfg3a", note="algorithmic bytes are the one-pass floor of SURVEY.md 8(d)")}
    if not args.no_e2e:
        def hstep():
            a = sl.compress(nr, nc, r32, c32, v_)           # 32-bit index arrays, as the C ABI takes them
            return a.pointers, a.indices, a.values
        dt, reps = timed_host(hstep)
        entry["e2e"] = {"value": n_in / dt, "unit": "triplets/s", "ms_per_step": dt * 1e3, "steps": len(reps), "ms_per_repeat": reps,
                        "h2d_bytes_per_step": int(n_in * 16), "d2h_bytes_per_step": int(nnz * 12 + (nc + 1) * 4),
                        "api": "slc_d_compress (pageable host triplets) + slc_d_download (fresh numpy buffers); median of the repeats"}
    if not args.no_cpu:
        t0 = time.perf_counter()
        want = oracle.compress(nr, nc, r_, c_, v_)
        dt = time.perf_counter() - t0
        entry["parity"] = {"oracle": "oracle.compress (Sparse.hs:184-291; duplicates summed in input order -- the reference's own "
                                     "order is unspecified, unstable introsort)", "bit_exact": same_matrix(built, want),
                           "equals_source_matrix_bit_exact": same_matrix(built, oracle.csc(*host_m)), "required": "structure bit-exact, values 1e-12"}
        entry["parity"]["ok"] = entry["parity"]["bit_exact"]
        entry["cpu_baseline"] = {"value": n_in / dt, "unit": "triplets/s", "cores": 1, "kind": "port",
                                 "sample": "all %d triplets, one pass: %.2f s" % (n_in, dt), "host_cores_available": os.cpu_count()}
    return entry


# ---- cfg 4: R-MAT generated on the device ---------------------------------------------------------------

def rmat_device(scale, edge_factor=16, seed=0, rows=None):
    """R-MAT triplets (a, b, c, d) = (0.57, 0.19, 0.19, 0.05) generated with torch on the GPU (numpy needs
    ~95 s for scale 24 on one core; this takes a fraction of a second).  Deterministic for a given seed,
    GPU model and torch version: every rank of a multi-GPU run generates the same list."""
    import torch
    gen = torch.Generator(device="cuda")
    gen.manual_seed(1000 + seed)
    n = 1 << scale
    mm = n * edge_factor
    r = torch.zeros(mm, dtype=torch.int32, device="cuda")
    c = torch.zeros(mm, dtype=torch.int32, device="cuda")
    a, b, cc = 0.57, 0.19, 0.19
    for bit in range(scale):
        u = torch.rand(mm, dtype=torch.float32, device="cuda", generator=gen)
        r |= (u >= (a + b)).to(torch.int32) << bit
        c |= (((u >= a) & (u < a + b)) | (u >= a + b + cc)).to(torch.int32) << bit
        del u
    v = torch.rand(mm, dtype=torch.float64, device="cuda", generator=gen) * 2 - 1
    return n, r, c, v


def cfg4_build(sl, scale):
    import torch
    n, r, c, v = rmat_device(scale)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    m = sl.compress_dev(n, n, r.numel(), r.data_ptr(), c.data_ptr(), v.data_ptr(), np.float64)
    sl.synchronize()
    asm_ms = (time.perf_counter() - t0) * 1e3
    info = {"triplets": int(r.numel()), "nnz_after_duplicate_merge": sl.nonZero(m), "device_assembly_ms": asm_ms,
            "triplets_per_s": r.numel() / (asm_ms * 1e-3)}
    del r, c, v
    torch.cuda.empty_cache()
    return m, info


# ---- the reference arm ------------------------------------------------------------------------------------

def run_reference(args, rank):
    """--impl reference: the reference's CPU algorithm (oracle port; the Haskell original cannot be built
    here) on the workload --workload names, rank 0 only.  Each step is a bounded sample of that workload
    (a column block for the products, a smaller grid for the structural ops) so K + W steps end within
    minutes; the metric is a rate, so the sample does not change its meaning."""
    if rank != 0:
        return
    import oracle
    from sparse_linear_b200 import generators as g
    w = args.workload
    note = ("oracle/ C port of the cited Haskell (single-threaded by construction, SURVEY.md section 0.2); no GHC exists in "
            "this image so the reference itself cannot be built")
    unit, metric = "GB/s", METRIC
    if w in ("cfg5", "cfg1", "cfg4"):
        if w == "cfg4":
            scale = args.grid or 20                       # the numpy generator needs ~6 s per 16 M edges
            n, r, c, v = g.rmat_triplets(scale, 16, seed=0)
            m = oracle.compress(n, n, r, c, v)
            desc = "cfg4: R-MAT scale %d edge factor 16, Double SpMV (sample: scale 24 takes minutes to generate on one core)" % scale
        else:
            m = g.laplace2d_5pt(args.grid or 1024) if w == "cfg1" else g.stencil3d_27pt(args.grid or 200)
            desc = ("cfg1: 2D 5-point Laplacian, Double SpMV (mulV)" if w == "cfg1" else
                    "cfg5: 3D 27-point stencil %d^3, Double SpMV (mulV)" % (args.grid or 200))
        frac_cols = max(1, m.ncols // 8) if m.colptr[-1] > 50_000_000 else m.ncols
        nnz = int(m.colptr[frac_cols])
        sub = oracle.csc(m.nrows, frac_cols, m.colptr[:frac_cols + 1], m.rowidx[:nnz], m.values[:nnz])
        x = g.uniform_vector(frac_cols, seed=0)
        units = nnz * 12 + (frac_cols + 1) * 4 + frac_cols * 8 + m.nrows * 8
        sample = "columns [0,%d) of %d (%d nnz) per step" % (frac_cols, m.ncols, nnz)

        def step():
            oracle.mulv(sub, x)
        scale_units = 1e9
    elif w in ("cfg2", "cfg5mm"):
        cx = w == "cfg2"
        m = g.laplace2d_5pt(args.grid or 1024, np.complex128) if cx else g.stencil3d_27pt(args.grid or 100)
        val = 16 if cx else 8
        k, kk = 64, 2
        om = oracle.csc(*m)
        X = g.uniform_block(m.ncols, kk, seed=0, dtype=np.complex128 if cx else np.float64)
        units = g.spmm_bytes(m.nrows, m.ncols, int(m.colptr[-1]), k, val) * kk / k
        desc = ("cfg2: 2D 5-point Laplacian 1024^2, Complex Double SpMM k=64" if cx else
                "cfg5: 3D 27-point stencil %d^3 (sample grid), Double SpMM k=64" % (args.grid or 100))
        sample = "%d of 64 columns per step (the reference serves the columns one by one, Feast.hs:203-208)" % kk
        metric = "spmm_effective_hbm_gbps"

        def step():
            oracle.mulm(om, X)
        scale_units = 1e9
    elif w == "cfg3":
        n = args.grid or 64
        m = g.laplace3d_7pt(n)
        om = oracle.csc(*m)
        units = int(oracle.mm(om, om).colptr[-1])
        desc = "cfg3: SpGEMM A*A, 3D 7-point Laplacian (fast restatement: output-identical touched-list accumulator)"
        sample = "%d^3 grid per step (%d output entries)" % (n, units)
        unit, metric, scale_units = "nnz/s", "spgemm_output_nnz_per_s", 1.0

        def step():
            oracle.mm(om, om)
    else:
        n = args.grid or 64
        m = g.laplace3d_7pt(n) if w.startswith("cfg3") else g.stencil3d_27pt(args.grid or 64)
        om = oracle.csc(*m)
        if w.endswith("a"):
            r, c, v = g.split_duplicates(m, seed=0)
            units = len(r)
            unit, metric = "triplets/s", "assembly_triplets_per_s"
            desc = "cfg3: triplet assembly with 2x duplicates"

            def step():
                oracle.compress(m.nrows, m.ncols, r, c, v)
        else:
            units = int(m.colptr[-1])
            unit, metric = "nnz/s", "transpose_nnz_per_s"
            desc = "%s: transpose" % w[:4]

            def step():
                oracle.transpose(om)
        sample = "%d^3 grid per step" % n
        scale_units = 1.0
    for _ in range(min(args.warmup, 3)):
        step()
    steps = args.steps
    t0 = time.perf_counter()
    done = 0
    for _ in range(steps):
        step()
        done += 1
        if time.perf_counter() - t0 > 150.0:     # bound the arm to a few minutes whatever K is
            break
    dt = time.perf_counter() - t0
    value = units * done / dt / scale_units
    line = {
        "impl": "reference", "metric": metric, "value": value, "unit": unit, "n_gpus": args.gpus,
        "steps": done, "warmup": min(args.warmup, 3), "ms_per_step": dt / done * 1e3,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "c128" if w == "cfg2" else "f64",
        "data": "synthetic (seeded)", "config": {"workload": desc, "sample": sample},
        "cpu_baseline": {"value": value, "unit": unit, "cores": 1, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": note,
    }
    print(json.dumps(line), flush=True)


# ---- ours ---------------------------------------------------------------------------------------------

def guarded(suite, name, fn):
    """One suite entry; a failure is recorded, never fatal to the headline line."""
    t0 = time.perf_counter()
    try:
        e = fn()
        if isinstance(e, tuple):
            e = e[0]
    except Exception as ex:  # noqa: BLE001
        e = {"name": name, "error": repr(ex), "trace": traceback.format_exc()[-1500:]}
    e["wall_s"] = round(time.perf_counter() - t0, 2)
    suite.append(e)
    return e


def line_from_entry(args, entry, suite=None):
    line = {"metric": entry.get("metric"), "value": entry.get("value"), "unit": entry.get("unit"), "n_gpus": 1,
            "steps": entry.get("steps"), "warmup": max(args.warmup, 3), "ms_per_step": entry.get("ms_per_step"),
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": entry.get("dtype"),
            "data": "synthetic (seeded numpy PCG64 / torch Philox)"}
    for k in ("config", "clocks", "gpu_launches", "e2e", "e2e_pageable", "roofline", "cpu_baseline", "parity", "cold_l2", "cuda_graph",
              "through_feast", "error", "trace"):
        if k in entry:
            line[k] = entry[k]
    if suite is not None:
        line["suite"] = suite
    return line


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
        import bench_multi
        return bench_multi.main(args, rank, world, local_rank, dist)

    sl.set_stream(torch.cuda.current_stream().cuda_stream)
    tm = Timer(sl, local_rank)
    w = args.workload
    n = args.grid

    def cfg5_spmv():
        q = n or 200
        hm = g.stencil3d_27pt(q)
        return entry_spmv(args, sl, tm, "cfg5_spmv", "cfg5: 3D 27-point stencil %d^3 (%d rows), Double SpMV (mulV)" % (q, q ** 3),
                          host_m=hm, traffic_key="cfg5") + (hm,)

    def cfg1():
        q = n or 1024
        return entry_spmv(args, sl, tm, "cfg1_spmv", "cfg1: 2D 5-point Laplacian %dx%d, Double SpMV (mulV); 84 MB working set: "
                          "fits L2, so warm, cold-L2 and CUDA-graph figures are reported separately" % (q, q),
                          host_m=g.laplace2d_5pt(q), graph_batch=True, cold=True, cpu_budget=1.0)

    def cfg2():
        q = n or 1024
        return entry_spmm(args, sl, tm, "cfg2_spmm", "cfg2: 2D 5-point Laplacian %dx%d, Complex Double SpMM k=64 (FEAST subspace block)" % (q, q),
                          g.laplace2d_5pt(q, np.complex128), np.complex128, 64, list(range(64)), traffic_key="cfg2", feast_leg=True)

    def cfg3():
        q = n or 128
        return entry_spgemm(args, sl, tm, g.laplace3d_7pt(q), q)

    def cfg3t():
        q = n or 128
        hm = g.laplace3d_7pt(q)
        m = sl.Matrix(hm.ncols, hm.nrows, hm.colptr, hm.rowidx, hm.values)
        return entry_transpose(args, sl, tm, "cfg3_transpose", "cfg3: transpose, 3D 7-point Laplacian %d^3 (handle -> handle)" % q, m, hm,
                               traffic_key="cfg3t")

    def cfg3a():
        q = n or 128
        return entry_assembly(args, sl, tm, g.laplace3d_7pt(q), q)

    def cfg4(which="both"):
        scale = n or 24
        m, info = cfg4_build(sl, scale)
        out = []
        if which in ("both", "spmv"):
            e, _ = entry_spmv(args, sl, tm, "cfg4_spmv", "cfg4: R-MAT scale %d edge factor 16 (generated and assembled on the device), "
                              "Double SpMV" % scale, dev_m=m, traffic_key="cfg4", cpu_budget=3.0)
            e["config"]["assembly"] = info
            out.append(e)
        if which in ("both", "t"):
            e = entry_transpose(args, sl, tm, "cfg4_transpose", "cfg4: R-MAT scale %d, transpose (handle -> handle)" % scale, m, None,
                                traffic_key="cfg4t", oracle_check=True)
            out.append(e)
        return out

    def cfg5mm(m=None, hm=None):
        q = n or 200
        hm = hm if hm is not None else g.stencil3d_27pt(q)
        return entry_spmm(args, sl, tm, "cfg5_spmm", "cfg5: 3D 27-point stencil %d^3, Double SpMM k=64" % q, hm, np.float64, 64,
                          [0, 21, 42, 63], traffic_key="cfg5mm", m=m)

    def cfg5t(m=None, hm=None):
        q = n or 200
        hm = hm if hm is not None else g.stencil3d_27pt(q)
        m = m if m is not None else sl.Matrix(hm.ncols, hm.nrows, hm.colptr, hm.rowidx, hm.values)
        return entry_transpose(args, sl, tm, "cfg5_transpose", "cfg5: transpose, 3D 27-point stencil %d^3" % q, m, hm,
                               traffic_key="cfg5t", oracle_check=True, symmetric=True)

    if w != "cfg5":            # a single workload as its own headline (development runs)
        single = {"cfg1": cfg1, "cfg2": cfg2, "cfg3": cfg3, "cfg3t": cfg3t, "cfg3a": cfg3a, "cfg5mm": cfg5mm, "cfg5t": cfg5t,
                  "cfg4": lambda: cfg4("spmv")[0], "cfg4t": lambda: cfg4("t")[0]}[w]
        e = single()
        if isinstance(e, tuple):
            e = e[0]
        print(json.dumps(line_from_entry(args, e)), flush=True)
        return

    head, m5, hm5 = cfg5_spmv()
    suite = []
    try:
        if args.suite == "auto" and not n:
            guarded(suite, "cfg5_spmm", lambda: cfg5mm(m5, hm5))
            guarded(suite, "cfg5_transpose", lambda: cfg5t(m5, hm5))
            del m5, hm5
            torch.cuda.empty_cache()
            guarded(suite, "cfg1_spmv", cfg1)
            guarded(suite, "cfg2_spmm", cfg2)
            guarded(suite, "cfg3_spgemm", cfg3)
            guarded(suite, "cfg3_transpose", cfg3t)
            guarded(suite, "cfg3_assembly", cfg3a)
            t0 = time.perf_counter()
            try:
                for e in cfg4():
                    e["wall_s"] = round(time.perf_counter() - t0, 2)
                    suite.append(e)
            except Exception as ex:  # noqa: BLE001
                suite.append({"name": "cfg4", "error": repr(ex), "trace": traceback.format_exc()[-1500:]})
    finally:
        line = line_from_entry(args, head, suite)
        line["steps"] = args.steps
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
