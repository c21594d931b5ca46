"""The PackedFunc registration shim (dgl-1_b200/csrc/packed.cu): libdgl_b200.so registers its
entry points into the registry of a loaded reference libdgl.so using only the reference's C ABI
(include/dgl/runtime/c_runtime_api.h:255-350), so the reference's Python would find them with
_init_api (python/dgl/_ffi/function.py:267-306).  Driven here through oracle/ref_ffi.py, i.e. the
same DGLFuncGetGlobal / DGLFuncCall calls the reference's ctypes FFI makes."""
import ctypes

import numpy as np
import pytest

from oracle import oracle, ref_ffi
from tests.util import random_graph

needs_ref = pytest.mark.skipif(not ref_ffi.available(), reason="oracle/_ref/libdgl.so not built")


def _register(override):
    from dgl_b200 import _lib
    lib = _lib.get_lib()
    n = lib.dglb200_register_packed_funcs(ctypes.c_void_p(ref_ffi.lib()._handle), int(override))
    assert n >= 5, _lib.last_error()
    return n


@needs_ref
def test_globals_registered_and_callable_without_gpu():
    assert _register(False) == 5
    assert ref_ffi.get_global("b200._CAPI_DGLB200Version")() == 1
    assert ref_ffi.get_global("b200._CAPI_DGLB200LaunchCount")() >= 0
    # errors travel through the reference's own convention: rc != 0 -> DGLGetLastError
    with pytest.raises(ref_ffi.RefError) as err:
        ref_ffi.get_global("b200._CAPI_DGLB200SpMMCsr")(1, 2, 3)
    assert "expects 16 arguments" in str(err.value)


def test_null_handle_is_an_error():
    from dgl_b200 import _lib
    assert _lib.get_lib().dglb200_register_packed_funcs(None, 0) == -1
    assert "NULL" in _lib.last_error()


@needs_ref
@pytest.mark.gpu
def test_degree_bucketing_override_matches_reference():
    """Call the reference's own global, then override it with the CUDA implementation and call the
    SAME name again: canonical forms must be bit-identical (the raw order of the reference is
    std::unordered_map iteration order)."""
    n = 500
    src, dst = random_graph(n, 6000, 5, hubs=3, hub_deg=400, zero_deg_frac=0.2)
    mids = np.arange(len(dst), dtype=np.int64)
    recv = np.unique(np.concatenate([dst, np.arange(n - 6, n)])).astype(np.int64)
    ref_raw = ref_ffi.degree_bucketing(mids, dst, recv)
    assert _register(True) == 6
    ours = ref_ffi.degree_bucketing(mids, dst, recv)          # now served by libdgl_b200.so
    for a, b in zip(oracle.canonicalize_bucketing(*ref_raw), ours):
        assert np.array_equal(a, b)
    for a, b in zip(ours, oracle.degree_bucketing(mids, dst, recv)):
        assert np.array_equal(a, b)
