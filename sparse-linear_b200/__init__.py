"""sparse-linear_b200: the B200-native compressed-sparse hot path of ttuegel/sparse-linear.

The product is libsparselinear_b200.so (hand-written sm_100a kernels behind the C ABI in
include/slc.h).  This package is the thin host layer tests and benchmarks drive: a ctypes
binding and a mirror of the reference's Data.Matrix.Sparse interface for the path.
Import as `sparse_linear_b200` (see sparse_linear_b200.py at the repo root).
"""
from . import _capi, feast, generators, sharded  # noqa: F401
from .feast import FeastWorkspace  # noqa: F401
from .sharded import ShardedMatrix, ShardedSpMV  # noqa: F401
from .matrix import (  # noqa: F401
    DeviceMatrix, Matrix, SparseLinearError,
    axpy, axpy_, cmap, compress, compress_dev, ctrans, decompress, device_count, diag, fromTriples, hermitian, ident,
    kronecker, launch_count, lin, mm, mulM, mulM_dev, mulV, mulV_dev, nonZero, scale, set_device,
    set_stream, spmm_window_error, synchronize, transpose, zeros,
)

__all__ = [n for n in dir() if not n.startswith("_")]
