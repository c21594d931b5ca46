"""ctypes driver for the UNMODIFIED reference C++ (oracle/_ref/libdgl.so, built by oracle/Makefile
`ref` straight from /root/reference/src).  TEST INFRASTRUCTURE ONLY.

Calls the reference's PackedFunc globals through its own C ABI (DGLFuncGetGlobal / DGLFuncCall,
include/dgl/runtime/c_runtime_api.h:255-362) without the reference's Python package, so it also
works on the GPU box where /root/reference does not exist but the prebuilt .so does:

  graph_index._CAPI_DGLGraphCreate / GetAdj / InDegrees   (src/graph/graph_apis.cc:105,324,474)
  runtime.degree_bucketing._CAPI_DGLDegreeBucketing       (src/scheduler/scheduler_apis.cc:16-23)

Arguments follow python/dgl/_ffi/_ctypes/function.py:81-139: ints as kDLInt int64, strings as
kStr, id arrays as kArrayHandle pointing at a DLTensor (int64, CPU, 1-D;
src/c_api_common.h:35-38); multi-value returns are PackedFuncs indexed by int
(src/c_api_common.cc:25-35).
"""
import ctypes
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_ref", "libdgl.so")

K_INT, K_HANDLE, K_NULL, K_ARRAY, K_FUNC, K_STR, K_NDARRAY = 0, 3, 4, 7, 10, 11, 13


class DLContext(ctypes.Structure):
    _fields_ = [("device_type", ctypes.c_int), ("device_id", ctypes.c_int)]


class DLDataType(ctypes.Structure):
    _fields_ = [("code", ctypes.c_uint8), ("bits", ctypes.c_uint8), ("lanes", ctypes.c_uint16)]


class DLTensor(ctypes.Structure):
    _fields_ = [("data", ctypes.c_void_p), ("ctx", DLContext), ("ndim", ctypes.c_int),
                ("dtype", DLDataType), ("shape", ctypes.POINTER(ctypes.c_int64)),
                ("strides", ctypes.POINTER(ctypes.c_int64)), ("byte_offset", ctypes.c_uint64)]


class DGLValue(ctypes.Union):
    _fields_ = [("v_int64", ctypes.c_int64), ("v_float64", ctypes.c_double),
                ("v_handle", ctypes.c_void_p), ("v_str", ctypes.c_char_p)]


def available():
    return os.path.exists(LIB_PATH)


_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        _LIB = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
        _LIB.DGLGetLastError.restype = ctypes.c_char_p
    return _LIB


class RefError(RuntimeError):
    pass


def _check(rc):
    if rc != 0:
        raise RefError(lib().DGLGetLastError().decode("utf-8", "replace"))


class _Func(object):
    """A PackedFunc handle (global or returned)."""

    def __init__(self, handle, is_global):
        self.handle = handle
        self.is_global = is_global

    def __del__(self):
        if not self.is_global and _LIB is not None:
            _LIB.DGLFuncFree(ctypes.c_void_p(self.handle))

    def __call__(self, *args):
        n = len(args)
        values = (DGLValue * max(n, 1))()
        codes = (ctypes.c_int * max(n, 1))()
        keep = []
        for i, a in enumerate(args):
            if isinstance(a, (bool, int, np.integer)):
                values[i].v_int64 = int(a)
                codes[i] = K_INT
            elif isinstance(a, str):
                b = a.encode()
                keep.append(b)
                values[i].v_str = b
                codes[i] = K_STR
            elif isinstance(a, np.ndarray):
                arr = np.ascontiguousarray(a, dtype=np.int64)
                shape = (ctypes.c_int64 * 1)(arr.shape[0])
                t = DLTensor(arr.ctypes.data, DLContext(1, 0), 1, DLDataType(0, 64, 1), shape, None, 0)
                keep += [arr, shape, t]
                values[i].v_handle = ctypes.cast(ctypes.pointer(t), ctypes.c_void_p)
                codes[i] = K_ARRAY
            elif isinstance(a, ctypes.c_void_p):
                values[i].v_handle = a
                codes[i] = K_HANDLE
            else:
                raise TypeError("unsupported argument %r" % (a,))
        ret = DGLValue()
        rcode = ctypes.c_int()
        _check(lib().DGLFuncCall(ctypes.c_void_p(self.handle), values, codes, n,
                                 ctypes.byref(ret), ctypes.byref(rcode)))
        rc = rcode.value
        if rc == K_INT:
            return ret.v_int64
        if rc == K_NULL:
            return None
        if rc == K_HANDLE:
            return ctypes.c_void_p(ret.v_handle)
        if rc == K_FUNC:
            return _Func(ret.v_handle, False)
        if rc in (K_NDARRAY, K_ARRAY):
            t = ctypes.cast(ret.v_handle, ctypes.POINTER(DLTensor)).contents
            n_el = 1
            for k in range(t.ndim):
                n_el *= t.shape[k]
            assert t.dtype.code == 0 and t.dtype.bits == 64
            buf = (ctypes.c_int64 * n_el).from_address(t.data + t.byte_offset) if n_el else []
            out = np.array(buf, dtype=np.int64).reshape(-1)
            if rc == K_NDARRAY:
                lib().DGLArrayFree(ctypes.c_void_p(ret.v_handle))   # DecRef the returned container
            return out
        raise RefError("unsupported return type code %d" % rc)


def get_global(name):
    h = ctypes.c_void_p()
    _check(lib().DGLFuncGetGlobal(name.encode(), ctypes.byref(h)))
    if not h.value:
        raise RefError("global %s is not registered" % name)
    return _Func(h.value, True)


class RefGraph(object):
    """The reference's Graph (mutable) or ImmutableGraph (readonly) built from an edge list."""

    def __init__(self, src, dst, n, readonly=False, multigraph=True):
        src = np.ascontiguousarray(src, dtype=np.int64)
        dst = np.ascontiguousarray(dst, dtype=np.int64)
        eid = np.arange(len(src), dtype=np.int64)
        self.handle = get_global("graph_index._CAPI_DGLGraphCreate")(
            src, dst, eid, bool(multigraph), int(n), bool(readonly))

    def __del__(self):
        try:
            get_global("graph_index._CAPI_DGLGraphFree")(self.handle)
        except Exception:  # noqa: BLE001
            pass

    def get_adj_csr(self, transpose=False):
        """Graph::GetAdj(transpose, "csr") -> (indptr, indices, eid), src/graph/graph.cc:525-549."""
        rst = get_global("graph_index._CAPI_DGLGraphGetAdj")(self.handle, bool(transpose), "csr")
        return rst(0), rst(1), rst(2)

    def get_adj_coo(self, transpose=False):
        rst = get_global("graph_index._CAPI_DGLGraphGetAdj")(self.handle, bool(transpose), "coo")
        return rst(0), rst(1)

    def in_degrees(self, v):
        return get_global("graph_index._CAPI_DGLGraphInDegrees")(self.handle, np.asarray(v, np.int64))


def _edge_array(name, handle, arg):
    rst = get_global(name)(handle, arg)
    return rst(0), rst(1), rst(2)


def graph_in_edges(g, v):
    """Graph::InEdges(IdArray) -> (src, dst, eid), src/graph/graph.cc:294-319."""
    return _edge_array("graph_index._CAPI_DGLGraphInEdges_2", g.handle, np.asarray(v, np.int64))


def graph_out_edges(g, v):
    return _edge_array("graph_index._CAPI_DGLGraphOutEdges_2", g.handle, np.asarray(v, np.int64))


def graph_edges(g, order=""):
    """Graph::Edges(order) -> (src, dst, eid), src/graph/graph.cc:368-409 ("srcdst" sorts)."""
    return _edge_array("graph_index._CAPI_DGLGraphEdges", g.handle, order)


def graph_edges_of_one(g, v, out=False):
    """Graph::InEdges(vid) / OutEdges(vid) -> (src, dst, eid), src/graph/graph.cc:284-292, 330-338."""
    rst = get_global("graph_index._CAPI_DGLGraph%sEdges_1" % ("Out" if out else "In"))(g.handle, int(v))
    return rst(0), rst(1), rst(2)


def graph_degree_of_one(g, v, out=False):
    """Graph::InDegree(vid) / OutDegree(vid), src/graph/graph.cc:402-409, 427-431."""
    return get_global("graph_index._CAPI_DGLGraph%sDegree" % ("Out" if out else "In"))(g.handle, int(v))


def graph_find_edges(g, eids):
    """Graph::FindEdges(eids) -> (src, dst, eid), src/graph/graph.cc:252-282."""
    return _edge_array("graph_index._CAPI_DGLGraphFindEdges", g.handle, np.asarray(eids, np.int64))


def graph_edge_ids(g, u, v):
    """Graph::EdgeIds(src, dst) -> (src, dst, eid): every edge between the pairs, src/graph/graph.cc:204-249."""
    rst = get_global("graph_index._CAPI_DGLGraphEdgeIds")(g.handle, np.asarray(u, np.int64), np.asarray(v, np.int64))
    return rst(0), rst(1), rst(2)


def graph_out_degrees(g, v):
    return get_global("graph_index._CAPI_DGLGraphOutDegrees")(g.handle, np.asarray(v, np.int64))


def graph_add_edges(g, src, dst):
    get_global("graph_index._CAPI_DGLGraphAddEdges")(g.handle, np.asarray(src, np.int64), np.asarray(dst, np.int64))


def degree_bucketing(msg_ids, vids, recv_ids):
    """sched::DegreeBucketing in the reference's own (unordered_map) order: 5 int64 arrays."""
    b = get_global("runtime.degree_bucketing._CAPI_DGLDegreeBucketing")(
        np.asarray(msg_ids, np.int64), np.asarray(vids, np.int64), np.asarray(recv_ids, np.int64))
    return tuple(b(i) for i in range(5))


def graph_uniform_sampling(g, seed_arrays, neighbor_type, num_hops, fanout, want_graph=False):
    """ImmutableGraph::NeighborUniformSample for 1, 2, 4 .. 128 seed arrays in ONE call (sampled on OpenMP
    threads, one per array: src/graph/graph_apis.cc:434-472).  `g` must be readonly.  Returns a list of
    (induced_nodes, induced_edges, layer_ids) int64 arrays; with want_graph also the subgraph's in-/out-CSR
    (indptr, indices, eid) as a 4th entry."""
    k = len(seed_arrays)
    assert k in (1, 2, 4, 8, 16, 32, 64, 128)
    name = "graph_index._CAPI_DGLGraphUniformSampling" + ("" if k == 1 else str(k))
    args = [np.ascontiguousarray(s, dtype=np.int64) for s in seed_arrays]
    rst = get_global(name)(g.handle, *args, neighbor_type, int(num_hops), int(fanout), k)
    out = []
    for i in range(k):
        handle = rst(i)
        rec = (rst(k + i), rst(2 * k + i), rst(3 * k + i))
        if want_graph:
            adj = get_global("graph_index._CAPI_DGLGraphGetAdj")(handle, neighbor_type != "in", "csr")
            rec = rec + ((adj(0), adj(1), adj(2)),)
        get_global("graph_index._CAPI_DGLGraphFree")(handle)
        out.append(rec)
    return out
