"""Run as a SUBPROCESS by tests/test_packed_registry.py (overriding globals of the loaded reference
libdgl.so is process-wide: isolation keeps the oracle of every other test unmodified).

  python tests/packed_graph_driver.py emu|gpu

1. asks the UNMODIFIED compiled reference (oracle/_ref/libdgl.so, through its own C ABI) for GetAdj csr/coo
   both ways, InEdges / OutEdges of a node list, Edges in both orders, In/OutDegrees;
2. registers libdgl_b200's overrides of the same global names (dglb200_register_packed_funcs, override=1);
3. makes the SAME calls again - now served by the device-resident index - and compares bit for bit;
4. mutates the graph through the reference's own AddEdges (wrapped: it must drop the cached index) and
   checks the answers against the numpy oracle of the mutated edge list; checks the error path.
"""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "dgl-1_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

from oracle import oracle, ref_ffi  # noqa: E402
from tests.util import random_graph  # noqa: E402


def load(mode):
    if mode == "emu":
        from tests.native.cuda_emu import build
        lib = ctypes.CDLL(build.build())
    else:
        from dgl_b200 import _lib
        lib = _lib.get_lib()
    lib.dglb200_register_packed_funcs.restype = ctypes.c_int
    lib.dglb200_register_packed_funcs.argtypes = [ctypes.c_void_p, ctypes.c_int]
    lib.dglb200_last_error.restype = ctypes.c_char_p
    lib.dglb200_launch_count.restype = ctypes.c_int64
    return lib


def queries(g, n, vids):
    out = {}
    for tr in (False, True):
        out["csr%d" % tr] = g.get_adj_csr(tr)
        out["coo%d" % tr] = g.get_adj_coo(tr)
    out["in_edges"] = ref_ffi.graph_in_edges(g, vids)
    out["out_edges"] = ref_ffi.graph_out_edges(g, vids)
    out["edges"] = ref_ffi.graph_edges(g, "")
    s, d, e = ref_ffi.graph_edges(g, "srcdst")
    order = np.lexsort((e, d, s))            # the reference's std::sort leaves parallel edges unordered
    out["edges_srcdst"] = (s[order], d[order], e[order])
    out["in_deg"] = (g.in_degrees(vids),)
    out["out_deg"] = (ref_ffi.graph_out_degrees(g, vids),)
    return out


def same(a, b, what):
    for k in a:
        for i, (x, y) in enumerate(zip(a[k], b[k])):
            assert np.array_equal(x, y), "%s: %s[%d] differs" % (what, k, i)


def main():
    mode = sys.argv[1]
    lib = load(mode)
    n = 300
    src, dst = random_graph(n, 4000, 7, hubs=3, hub_deg=500, zero_deg_frac=0.2)
    rng = np.random.RandomState(0)
    vids = np.concatenate([rng.randint(0, n, 60), [0, n - 1, 5, 5]]).astype(np.int64)    # repeats allowed
    g = ref_ffi.RefGraph(src, dst, n)
    tiny = ref_ffi.RefGraph(np.array([0, 1, 1], np.int64), np.array([1, 2, 2], np.int64), 4)
    tv = np.array([2, 3, 0], np.int64)
    want, want_tiny = queries(g, n, vids), queries(tiny, 4, tv)
    recv = np.unique(np.concatenate([dst, np.arange(n - 6, n)])).astype(np.int64)
    global bucket_in, bucket_ref
    bucket_in = (np.arange(len(dst), dtype=np.int64), dst, recv)
    bucket_ref = ref_ffi.degree_bucketing(*bucket_in)        # std::unordered_map order, canonicalised below

    l0 = lib.dglb200_launch_count()
    k = lib.dglb200_register_packed_funcs(ctypes.c_void_p(ref_ffi.lib()._handle), 1)
    assert k >= 17, (k, lib.dglb200_last_error())
    got, got_tiny = queries(g, n, vids), queries(tiny, 4, tv)
    assert lib.dglb200_launch_count() > l0, "the overrides launched no kernel"
    same(want, got, "300-node multigraph")
    same(want_tiny, got_tiny, "tiny graph")
    # sorted order is itself canonical now
    s, d, e = ref_ffi.graph_edges(g, "srcdst")
    order = np.lexsort((e, d, s))
    assert np.array_equal(order, np.arange(len(s)))

    # mutation through the reference's own (wrapped) AddEdges must invalidate the device index
    more_s, more_d = rng.randint(0, n, 50).astype(np.int64), rng.randint(0, n, 50).astype(np.int64)
    ref_ffi.graph_add_edges(g, more_s, more_d)
    s2, d2 = np.concatenate([src, more_s]), np.concatenate([dst, more_d])
    ip, ix, ei = g.get_adj_csr(False)
    w_ip, w_ix, w_ei = oracle.get_adj_csr(s2, d2, n)
    assert np.array_equal(ip, w_ip) and np.array_equal(ix, w_ix) and np.array_equal(ei, w_ei)
    assert np.array_equal(g.in_degrees(vids), np.bincount(d2, minlength=n)[vids])
    es, ed, ee = ref_ffi.graph_edges(g, "")
    assert np.array_equal(es, s2) and np.array_equal(ed, d2) and np.array_equal(ee, np.arange(len(s2)))

    # errors travel through the reference's convention
    for bad_call in (lambda: g.in_degrees(np.array([n + 5], np.int64)),
                     lambda: ref_ffi.graph_in_edges(g, np.array([-1], np.int64)),
                     lambda: ref_ffi.get_global("graph_index._CAPI_DGLGraphGetAdj")(g.handle, False, "xyz")):
        try:
            bad_call()
        except ref_ffi.RefError:
            pass
        else:
            raise AssertionError("an invalid query did not raise")
    del g, tiny            # _CAPI_DGLGraphFree (wrapped) drops the device index
    print("PACKED_GRAPH_OK")
    # degree bucketing: the override against the reference's raw output taken BEFORE the override
    ours = ref_ffi.degree_bucketing(*bucket_in)
    for a, b in zip(oracle.canonicalize_bucketing(*bucket_ref), ours):
        assert np.array_equal(a, b)
    for a, b in zip(ours, oracle.degree_bucketing(*bucket_in)):
        assert np.array_equal(a, b)
    print("BUCKETING_OK")


if __name__ == "__main__":
    main()
