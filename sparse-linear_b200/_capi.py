"""ctypes binding of libsparselinear_b200.so (include/slc.h).

There is no fallback: if the library has not been built, or no CUDA device is present when a
compute entry point is called, this raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libsparselinear_b200.so")

OK = 0
ERR_INVALID, ERR_DIM, ERR_ROW_BOUNDS, ERR_COL_BOUNDS = -1, -2, -3, -4
ERR_CUDA, ERR_NOMEM, ERR_FORMAT, ERR_RANGE = -5, -6, -7, -8

_i32p = C.POINTER(C.c_int32)
_f64p = C.POINTER(C.c_double)
_vp = C.c_void_p
_hp = C.c_void_p  # slc_matrix*

# name -> (restype, argtypes); every symbol include/slc.h declares
SIGNATURES = {
    "slc_version": (C.c_int, []),
    "slc_last_error": (C.c_char_p, []),
    "slc_device_count": (C.c_int, [C.POINTER(C.c_int)]),
    "slc_set_device": (C.c_int, [C.c_int]),
    "slc_set_stream": (C.c_int, [_vp]),
    "slc_synchronize": (C.c_int, []),
    "slc_launch_count": (C.c_int64, []),
    "slc_host_alloc": (C.c_int, [C.POINTER(_vp), C.c_int64]),
    "slc_host_free": (C.c_int, [_vp]),
    "slc_dev_alloc": (C.c_int, [C.POINTER(_vp), C.c_int64]),
    "slc_dev_free": (C.c_int, [_vp]),
    "slc_memcpy_h2d": (C.c_int, [_vp, _vp, C.c_int64]),
    "slc_memcpy_d2h": (C.c_int, [_vp, _vp, C.c_int64]),
    "slc_memcpy_d2d": (C.c_int, [_vp, _vp, C.c_int64]),
    "slc_free": (None, [_hp]),
    "slc_dims": (C.c_int, [_hp, _i32p, _i32p, C.POINTER(C.c_int64)]),
    "slc_elem_type": (C.c_int, [_hp]),
    "slc_prepare": (C.c_int, [_hp]),
    "slc_spmm_window_error": (C.c_int, [C.POINTER(C.c_int)]),
    "slc_set_spmv_kernel": (C.c_int, [_hp, C.c_int]),
    "slc_plan_info": (C.c_int, [_hp, C.c_char_p, C.c_int32]),
    "slc_ipc_export": (C.c_int, [_vp, _vp]),
    "slc_ipc_import": (C.c_int, [_vp, C.POINTER(_vp)]),
    "slc_ipc_close": (C.c_int, [_vp]),
    "slc_peer_group_create": (C.c_int, [C.c_int32, C.c_int32, C.POINTER(_vp), _vp]),
    "slc_peer_group_connect": (C.c_int, [_vp, _vp]),
    "slc_peer_barrier": (C.c_int, [_vp]),
    "slc_peer_barrier_all": (C.c_int, [_vp]),
    "slc_peer_signal": (C.c_int, [_vp]),
    "slc_peer_wait": (C.c_int, [_vp]),
    "slc_peer_group_timed_out": (C.c_int, [_vp, C.POINTER(C.c_int)]),
    "slc_peer_group_free": (None, [_vp]),
    "slc_peer_group_create_local": (C.c_int, [C.c_int32, _vp, C.POINTER(_vp)]),
    "slc_peer_group_set_deps": (C.c_int, [_vp, C.c_uint32]),
    "slc_peer_group_stats": (C.c_int, [_vp, C.POINTER(C.c_int64), C.c_int]),
    "slc_sharded_free": (None, [_vp]),
    "slc_sharded_dims": (C.c_int, [_vp, _i32p, _i32p, _i32p, C.POINTER(C.c_int64)]),
    "slc_sharded_splits": (C.c_int, [_vp, _vp, _vp]),
    "slc_sharded_block": (C.c_int, [_vp, C.c_int32, C.POINTER(_hp), _i32p, C.POINTER(_vp)]),
    "slc_sharded_operand": (C.c_int, [_vp, C.c_int32, C.c_int32, C.POINTER(_vp), C.POINTER(C.c_int64)]),
    "slc_sharded_result": (C.c_int, [_vp, C.c_int32, C.POINTER(_vp), C.POINTER(C.c_int64)]),
    "slc_sharded_synchronize": (C.c_int, [_vp]),
    "slc_device_arrays": (C.c_int, [_hp, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp)]),
    "slc_coo_bucket_noalloc_dev": (C.c_int, [C.c_int, C.c_int64, _vp, _vp, _vp, C.c_int32, _vp, _vp, _vp, _vp, _vp]),
}
for _p in ("d", "z"):
    SIGNATURES.update({
        "slc_%s_upload" % _p: (C.c_int, [C.c_int32, C.c_int32, _vp, _vp, _vp, C.POINTER(_hp)]),
        "slc_%s_download" % _p: (C.c_int, [_hp, _vp, _vp, _vp]),
        "slc_%s_mulv" % _p: (C.c_int, [_hp, _vp, _vp]),
        "slc_%s_axpy" % _p: (C.c_int, [_hp, _vp, _vp]),
        "slc_%s_mulv_dev" % _p: (C.c_int, [_hp, _vp, _vp]),
        "slc_%s_axpy_dev" % _p: (C.c_int, [_hp, _vp, _vp]),
        "slc_%s_mulv_dist_dev" % _p: (C.c_int, [_hp, C.c_int32, C.c_int32, _vp, _vp, _vp, _vp]),
        "slc_%s_mulv_allgather_dev" % _p: (C.c_int, [_hp, C.c_int32, C.c_int32, _vp, _vp, _vp, _vp, _vp]),
        "slc_%s_mulm_dist_dev" % _p: (C.c_int, [_hp, C.c_int32, C.c_int32, _vp, _vp, _vp, _vp, C.c_int64,
                                                C.c_int32, _vp]),
        "slc_%s_mulm" % _p: (C.c_int, [_hp, _vp, C.c_int64, _vp, C.c_int64, C.c_int32, _vp]),
        "slc_%s_mulm_dev" % _p: (C.c_int, [_hp, _vp, C.c_int64, _vp, C.c_int64, C.c_int32, _vp]),
        "slc_%s_dist_needs" % _p: (C.c_int, [_hp, C.c_int32, C.c_int32, _vp, C.POINTER(C.c_uint32)]),
        "slc_%s_upload_sharded" % _p: (C.c_int, [C.c_int32, _vp, C.c_int32, C.c_int32, _vp, _vp, _vp, C.POINTER(_vp)]),
        "slc_%s_mulv_sharded" % _p: (C.c_int, [_vp, _vp, _vp]),
        "slc_%s_mulm_sharded" % _p: (C.c_int, [_vp, _vp, C.c_int64, _vp, C.c_int64, C.c_int32]),
        "slc_%s_vstack_dev" % _p: (C.c_int, [C.c_int32, _vp, C.c_int32, _vp, _vp, _vp, C.POINTER(_hp)]),
        "slc_%s_coo_bucket_dev" % _p: (C.c_int, [C.c_int64, _vp, _vp, _vp, C.c_int32, _vp, _vp, _vp, _vp, _vp]),
        "slc_%s_mulv_sharded_dev" % _p: (C.c_int, [_vp]),
        "slc_%s_mulm_sharded_dev" % _p: (C.c_int, [_vp, C.c_int32]),
        "slc_%s_mm" % _p: (C.c_int, [_hp, _hp, C.POINTER(_hp)]),
        "slc_%s_transpose" % _p: (C.c_int, [_hp, C.c_int, C.POINTER(_hp)]),
        "slc_%s_hermitian" % _p: (C.c_int, [_hp, C.POINTER(C.c_int)]),
        "slc_%s_compress" % _p: (C.c_int, [C.c_int32, C.c_int32, C.c_int64, _vp, _vp, _vp,
                                           C.POINTER(_hp), C.POINTER(C.c_int64)]),
        "slc_%s_compress_dev" % _p: (C.c_int, [C.c_int32, C.c_int32, C.c_int64, _vp, _vp, _vp,
                                               C.POINTER(_hp), C.POINTER(C.c_int64)]),
        "slc_%s_lin" % _p: (C.c_int, [_vp, _hp, _vp, _hp, C.POINTER(_hp)]),
        "slc_%s_kronecker" % _p: (C.c_int, [_hp, _hp, C.POINTER(_hp)]),
    })

_lib = None


class LibraryMissing(RuntimeError):
    pass


def lib():
    """Load the CUDA library (once).  Fails loudly when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise LibraryMissing(
                "%s is missing: run `python sparse-linear_b200/build.py` (or __graft_entry__.build()); "
                "there is no CPU fallback" % LIB_PATH)
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError here = header/library mismatch
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def last_error():
    return lib().slc_last_error().decode("utf-8", "replace")
