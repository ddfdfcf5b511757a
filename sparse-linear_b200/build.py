"""Build libsparselinear_b200.so in-tree with nvcc for sm_100a.

Usage:  python sparse-linear_b200/build.py [--force] [--verbose]

Each .cu under csrc/ is compiled to build/*.o (in parallel, skipped when up to date) and
linked into sparse-linear_b200/libsparselinear_b200.so.  cudart is linked statically, so
the library only needs the NVIDIA driver at run time.
"""
import concurrent.futures
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
OBJDIR = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libsparselinear_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")

FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC,-fvisibility=hidden",
    "-I", os.path.join(ROOT, "include"), "-I", CSRC,
    "--expt-relaxed-constexpr",
]


def _newest(paths):
    return max(os.path.getmtime(p) for p in paths)


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def headers():
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    hs.append(os.path.join(ROOT, "include", "slc.h"))
    return hs


def _compile(src, verbose, extra):
    obj = os.path.join(OBJDIR, os.path.basename(src)[:-3] + ".o")
    cmd = [NVCC, *FLAGS, *extra, "-c", src, "-o", obj]
    if verbose:
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    return src, obj, r.returncode, r.stdout + r.stderr


def build(force=False, verbose=False, ptxas_verbose=False):
    os.makedirs(OBJDIR, exist_ok=True)
    srcs = sources()
    hdr_time = _newest(headers() + [os.path.abspath(__file__)])
    todo, objs = [], []
    for s in srcs:
        obj = os.path.join(OBJDIR, os.path.basename(s)[:-3] + ".o")
        objs.append(obj)
        if force or not os.path.exists(obj) or os.path.getmtime(obj) < max(os.path.getmtime(s), hdr_time):
            todo.append(s)
    extra = ["-Xptxas", "-v"] if ptxas_verbose else []
    if todo:
        with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, len(todo))) as ex:
            for src, obj, rc, out in ex.map(lambda s: _compile(s, verbose, extra), todo):
                if out.strip() and (rc != 0 or verbose or ptxas_verbose):
                    print(out, flush=True)
                if rc != 0:
                    raise RuntimeError("nvcc failed on %s" % src)
    if todo or force or not os.path.exists(LIB) or os.path.getmtime(LIB) < _newest(objs):
        cmd = [NVCC, "-shared", "-o", LIB, *objs, "-cudart", "static"]
        if verbose:
            print(" ".join(cmd), flush=True)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            print(r.stdout + r.stderr)
            raise RuntimeError("link failed")
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv,
                ptxas_verbose="--ptxas" in sys.argv))
