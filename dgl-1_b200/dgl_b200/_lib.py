"""ctypes binding of libdgl_b200.so (the C ABI declared in include/dgl_b200.h).

Plays the role of python/dgl/_ffi/base.py:_load_lib + check_call in the reference
(python/dgl/_ffi/base.py:31-62): load the shared library, turn a non-zero return code into
DGLError carrying the library's thread-local message.  There is NO fallback: if the library is
missing or was not built, importing an op raises.
"""
import ctypes
import os

from .base import DGLError

_LIB_NAME = "libdgl_b200.so"

c_i64 = ctypes.c_int64
c_i32 = ctypes.c_int32
c_int = ctypes.c_int
c_vp = ctypes.c_void_p
c_sz = ctypes.c_size_t


class SpmmPlanStruct(ctypes.Structure):
    """Mirror of dglb200_spmm_plan_t (include/dgl_b200.h)."""
    _fields_ = [
        ("chunk", c_i32),
        ("n_long_rows", c_i32),
        ("n_chunks", c_i64),
        ("long_rows", c_vp),
        ("chunk_ptr", c_vp),
        ("partial", c_vp),
        ("partial_arg", c_vp),
        ("partial_bytes", c_sz),
    ]


class SpmmExtStruct(ctypes.Structure):
    """Mirror of dglb200_spmm_ext_t (include/dgl_b200.h)."""
    _fields_ = [("accumulate", c_i32), ("relu", c_i32), ("x_halo", c_vp), ("halo_first", c_i64),
                ("row_scale", c_vp), ("bias", c_vp)]


# name -> (restype, argtypes); must list every symbol of include/dgl_b200.h
class SampledSubgraphStruct(ctypes.Structure):
    """dglb200_sampled_subgraph_t (include/dgl_b200.h)."""
    _fields_ = [("n_nodes", ctypes.c_int64), ("n_edges", ctypes.c_int64), ("nodes", ctypes.c_void_p),
                ("layers", ctypes.c_void_p), ("sub_indptr", ctypes.c_void_p), ("sub_indices", ctypes.c_void_p),
                ("sub_rows", ctypes.c_void_p), ("induced_edges", ctypes.c_void_p), ("needed_bytes", ctypes.c_size_t)]


SIGNATURES = {
    "dglb200_abi_version": (c_int, []),
    "dglb200_last_error": (ctypes.c_char_p, []),
    "dglb200_launch_count": (c_i64, []),
    "dglb200_device_cc": (c_int, []),
    "dglb200_set_tuning": (c_int, [ctypes.c_char_p, c_i64]),
    "dglb200_csr_build_workspace_bytes": (c_int, [c_i64, c_i64, ctypes.POINTER(c_sz)]),
    "dglb200_csr_build": (c_int, [c_i64, c_i64, c_i64, c_vp, c_vp, c_vp, c_int, c_vp, c_vp, c_vp,
                                  c_vp, c_sz, c_vp]),
    "dglb200_widen_i32_to_i64": (c_int, [c_i64, c_vp, c_vp, c_vp]),
    "dglb200_degrees_from_indptr": (c_int, [c_i64, c_vp, c_vp, c_vp]),
    "dglb200_csr_row_ids": (c_int, [c_i64, c_i64, c_vp, c_vp, c_vp]),
    "dglb200_csr_select_workspace_bytes": (c_int, [c_i64, ctypes.POINTER(c_sz)]),
    "dglb200_csr_select_offsets": (c_int, [c_i64, c_vp, c_i64, c_vp, c_vp, ctypes.POINTER(c_i64), c_vp, c_sz, c_vp]),
    "dglb200_csr_select_expand": (c_int, [c_i64, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp,
                                          c_vp, c_vp]),
    "dglb200_degrees_of": (c_int, [c_i64, c_vp, c_i64, c_vp, c_vp, c_vp, c_vp]),
    "dglb200_neighbor_sample_offsets": (c_int, [c_i64, c_vp, c_vp, c_i64, c_vp, c_i64, c_vp, ctypes.POINTER(c_i64),
                                                ctypes.POINTER(c_i64), c_vp, c_sz, c_vp]),
    "dglb200_neighbor_sample_fill": (c_int, [c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_i64, ctypes.c_uint64,
                                             c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_i32, c_vp, c_vp, c_vp]),
    "dglb200_sample_subgraph": (c_int, [c_i64, c_vp, c_vp, c_vp, c_i64, c_vp, c_i64, c_i32, ctypes.c_uint64, c_vp,
                                        c_vp, c_vp, c_vp, c_sz, ctypes.POINTER(SampledSubgraphStruct), c_vp]),
    "dglb200_spmm_plan_workspace_bytes": (c_int, [c_i64, ctypes.POINTER(c_sz)]),
    "dglb200_spmm_plan": (c_int, [c_i64, c_vp, c_i32, c_i64, c_vp, c_vp, ctypes.POINTER(c_i64),
                                  c_vp, c_sz, c_vp]),
    "dglb200_spmm_csr": (c_int, [c_i64, c_i64, c_i64, c_vp, c_vp, c_vp, c_int, c_int, c_int,
                                 c_vp, c_i64, c_vp, c_i64, c_vp, c_vp,
                                 ctypes.POINTER(SpmmPlanStruct), c_vp]),
    "dglb200_spmm_csr_strided": (c_int, [c_i64, c_i64, c_i64, c_vp, c_vp, c_vp, c_int, c_int, c_int,
                                         c_vp, c_i64, c_i64, c_vp, c_i64, c_vp, c_i64, c_vp,
                                         ctypes.POINTER(SpmmPlanStruct), c_vp]),
    "dglb200_spmm_csr_ext": (c_int, [c_i64, c_i64, c_i64, c_vp, c_vp, c_vp, c_int, c_int, c_int,
                                     c_vp, c_i64, c_i64, c_vp, c_i64, c_vp, c_i64, c_vp,
                                     ctypes.POINTER(SpmmPlanStruct), ctypes.POINTER(SpmmExtStruct), c_vp]),
    "dglb200_spmm_csr_static_scratch": (c_int, [c_i64, c_i64, c_i64, c_vp, c_vp, c_vp, c_int, c_int, c_int,
                                                c_vp, c_i64, c_vp, c_i64, c_vp,
                                                ctypes.POINTER(SpmmPlanStruct), c_vp]),
    "dglb200_edge_data_to_slots": (c_int, [c_i64, c_vp, c_i64, c_i64, c_vp, c_vp, c_vp, c_vp]),
    "dglb200_spmm_csr_backward_x": (c_int, [c_i64, c_i64, c_i64, c_vp, c_vp, c_vp, c_int, c_vp, c_i64,
                                            c_vp, c_i64, c_vp, c_vp,
                                            ctypes.POINTER(SpmmPlanStruct), c_vp]),
    "dglb200_sddmm_csr": (c_int, [c_i64, c_i64, c_vp, c_vp, c_vp, c_int, c_int, c_vp, c_vp,
                                  c_i64, c_i64, c_vp, c_vp]),
    "dglb200_edge_softmax_csr": (c_int, [c_i64, c_i64, c_vp, c_vp, c_i64, c_vp, c_vp, c_vp]),
    "dglb200_edge_softmax_backward_csr": (c_int, [c_i64, c_i64, c_vp, c_vp, c_i64, c_vp, c_vp,
                                                  c_vp, c_vp]),
    "dglb200_gat_aggregate_csr": (c_int, [c_i64, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp,
                                          c_i64, c_i64, ctypes.c_float, c_int, c_vp, c_vp, c_vp,
                                          c_vp]),
    "dglb200_bucket_capacity": (c_i64, [c_i64]),
    "dglb200_degree_bucketing_workspace_bytes": (c_int, [c_i64, c_i64, ctypes.POINTER(c_sz)]),
    "dglb200_degree_bucketing": (c_int, [c_i64, c_i64, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp,
                                         c_vp, c_vp, ctypes.POINTER(c_i64), c_vp, c_sz, c_vp]),
    "dglb200_unique_workspace_bytes": (c_int, [c_i64, ctypes.POINTER(c_sz)]),
    "dglb200_unique_sorted": (c_int, [c_i64, c_i64, c_vp, c_vp, c_vp, ctypes.POINTER(c_i64),
                                      c_vp, c_sz, c_vp]),
    "dglb200_gather_rows": (c_int, [c_i64, c_vp, c_i64, c_vp, c_i64, c_vp, c_vp, c_vp]),
    "dglb200_gather_rows_strided": (c_int, [c_i64, c_vp, c_i64, c_vp, c_i64, c_i64, c_vp, c_i64, c_vp,
                                            c_int, c_vp]),
    "dglb200_gather_rows_convert": (c_int, [c_i64, c_vp, c_i64, c_int, c_vp, c_i64, c_i64, c_vp, c_i64, c_vp,
                                            c_int, c_vp]),
    "dglb200_gather_rows_multi": (c_int, [c_int, c_vp, c_vp, c_vp, c_vp, c_i64, c_vp, c_vp, c_int, c_vp]),
    "dglb200_ipc_get_handle": (c_int, [c_vp, c_vp, ctypes.POINTER(c_i64)]),
    "dglb200_ipc_open": (c_int, [c_vp, ctypes.POINTER(c_vp)]),
    "dglb200_ipc_close": (c_int, [c_vp]),
    "dglb200_peer_push": (c_int, [c_int, c_vp, c_vp, c_vp, c_vp, c_i64, c_vp, c_i64, c_i64, c_i64, c_int,
                                  c_vp, c_int, c_i32, c_i32, c_vp, c_vp, c_vp, c_int, c_int, c_vp]),
    "dglb200_peer_signal": (c_int, [c_int, c_vp, c_i32, c_vp]),
    "dglb200_peer_wait": (c_int, [c_vp, c_int, c_i32, c_vp, c_vp]),
    "dglb200_register_packed_funcs": (c_int, [c_vp, c_int]),
}


def lib_path():
    """Path of the in-tree shared library (override with DGL_B200_LIB, like DGL_LIBRARY_PATH
    in python/dgl/_ffi/libinfo.py:32-33)."""
    env = os.environ.get("DGL_B200_LIB")
    if env:
        return env
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib", _LIB_NAME)


_LIB = None


def get_lib():
    """Load (once) and return the ctypes handle; raises DGLError if the library is absent."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = lib_path()
    if not os.path.exists(path):
        raise DGLError(
            "%s not found at %s: build it with `python -c 'import __graft_entry__ as g; g.build()'`"
            " (make -C dgl-1_b200). There is no CPU fallback." % (_LIB_NAME, path))
    try:
        lib = ctypes.CDLL(path, mode=ctypes.RTLD_GLOBAL)
    except OSError as err:
        raise DGLError("cannot load %s: %s" % (path, err))
    for name, (res, args) in SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError:
            raise DGLError("%s does not export %s (stale build?)" % (path, name))
        fn.restype = res
        fn.argtypes = args
    if lib.dglb200_abi_version() != 1:
        raise DGLError("ABI version mismatch: library %d, binding 1" % lib.dglb200_abi_version())
    _LIB = lib
    return lib


def last_error():
    return get_lib().dglb200_last_error().decode("utf-8", "replace")


def check_call(rc):
    """Raise DGLError with the library's message on a non-zero return (cf. _ffi/base.py:50-62)."""
    if rc != 0:
        raise DGLError(last_error())


def launch_count():
    return int(get_lib().dglb200_launch_count())
