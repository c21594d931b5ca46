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
    assert _register(False) == 5          # override_reference = 0: only the b200.* names
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


def _run_driver(mode):
    """Overrides are process-wide in the loaded libdgl.so: the override tests run in a subprocess so that
    the reference stays unmodified for every other test of this process."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    p = subprocess.run([sys.executable, os.path.join(root, "tests", "packed_graph_driver.py"), mode],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=900)
    out = p.stdout.decode("utf-8", "replace")
    assert p.returncode == 0 and "PACKED_GRAPH_OK" in out and "BUCKETING_OK" in out, out[-3000:]
    assert "PACKED_SAMPLING_OK" in out, out[-3000:]


@needs_ref
def test_reference_globals_overridden_by_emulated_kernels():
    """graph_index._CAPI_DGLGraph{GetAdj, InEdges_2, OutEdges_2, Edges, InDegrees, OutDegrees} and
    runtime.degree_bucketing._CAPI_DGLDegreeBucketing of the compiled reference, overridden by
    libdgl_b200's PackedFuncs (same names, arity and return order) and called through DGLFuncCall like the
    reference's Python does: bit-identical to the unmodified reference; mutation through the reference's
    AddEdges invalidates the device index.  CPU tier: the kernels are the emulated CUDA sources."""
    import shutil
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    _run_driver("emu")


@needs_ref
@pytest.mark.gpu
def test_reference_globals_overridden_on_the_gpu():
    """The same, served by libdgl_b200.so on the B200."""
    _run_driver("gpu")
