"""The C-ABI library loads and exports every symbol include/slc.h declares (no compute calls:
this runs without a GPU)."""
import os
import re

import pytest

import sparse_linear_b200 as sl

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "slc.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(slc_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = sl._capi.lib()
    names = declared_symbols()
    assert len(names) >= 45
    for n in names:
        assert hasattr(lib, n), "library does not export %s" % n
    assert sorted(sl._capi.SIGNATURES) == names, "ctypes table and header disagree"


def test_version_and_error_string():
    lib = sl._capi.lib()
    assert lib.slc_version() >= 100
    assert isinstance(sl._capi.last_error(), str)


def test_no_cpu_fallback_without_device():
    if sl.device_count() > 0:
        pytest.skip("a CUDA device is present")
    import numpy as np
    with pytest.raises(sl.SparseLinearError):
        sl.mulV(sl.ident(4), np.ones(4))


def test_host_side_argument_checks_need_no_device():
    import numpy as np
    a = sl.ident(3)
    with pytest.raises(sl.SparseLinearError, match="axpy_: column dimension 3 does not match operand dimension 4"):
        sl.mulV(a, np.ones(4))
    with pytest.raises(sl.SparseLinearError, match="mm: inner dimension mismatch"):
        sl.mm(a, sl.ident(4))
    with pytest.raises(sl.SparseLinearError, match="glin: row number mismatch"):
        sl.lin(1, a, 1, sl.ident(4))
    with pytest.raises(sl.SparseLinearError, match="compress: row and column array lengths differ"):
        sl.compress(3, 3, [0, 1], [0], [1.0, 2.0])
