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


def queries(g, n, vids, readonly=False):
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
    ne = len(out["edges"][0])
    pick = np.array([0, ne - 1, ne // 2, ne // 2, 3 % ne], np.int64)
    es, ed = out["edges"][0], out["edges"][1]
    if not readonly:                      # FindEdges is mutable-graph only (immutable_graph.h:270); Edges("") is in
        out["find_edges"] = ref_ffi.graph_find_edges(g, pick)                          # edge-id order only there
        out["edge_ids_pairs"] = ref_ffi.graph_edge_ids(g, es[pick], ed[pick])          # incl. parallel edges
        out["edge_ids_bcast_src"] = ref_ffi.graph_edge_ids(g, es[pick[:1]], vids)
        out["edge_ids_bcast_dst"] = ref_ffi.graph_edge_ids(g, vids, ed[pick[:1]])
    for v in (int(vids[0]), int(vids[-1]), n - 1):          # single-vertex forms (ids that repeat, a zero-degree tail node)
        out["in_edges_1_%d" % v] = ref_ffi.graph_edges_of_one(g, v)
        out["out_edges_1_%d" % v] = ref_ffi.graph_edges_of_one(g, v, out=True)
        out["degree_1_%d" % v] = (np.array([ref_ffi.graph_degree_of_one(g, v), ref_ffi.graph_degree_of_one(g, v, out=True)]),)
    out["in_deg"] = (g.in_degrees(vids),)
    out["out_deg"] = (ref_ffi.graph_out_degrees(g, vids),)
    return out


def same(a, b, what):
    for k in a:
        for i, (x, y) in enumerate(zip(a[k], b[k])):
            assert np.array_equal(x, y), "%s: %s[%d] differs" % (what, k, i)


def sampling_cases(n, rng):
    """(seed arrays, neighbour type, fanout, hops): deterministic ones (fanout covers every degree) and random ones."""
    out = []
    for ntype in ("in", "out"):
        for fanout, hops, k in ((10 ** 6, 2, 1), (10 ** 6, 1, 4), (4, 2, 2), (1, 3, 1), (600, 1, 1)):
            seeds = [rng.randint(0, n, 25).astype(np.int64) for _ in range(k)]
            seeds[0][0] = 3                  # a hub
            out.append((seeds, ntype, fanout, hops))
    return out


def check_sampling(g, ro_src, ro_dst, n, cases, want):
    """Every case through graph_index._CAPI_DGLGraphUniformSampling*: the invariants of SampleSubgraph on the
    parent's immutable CSR; deterministic cases bit for bit against `want` (the unmodified reference's answers,
    taken before the override; None = just the invariants)."""
    for ci, (seeds, ntype, fanout, hops) in enumerate(cases):
        csr = oracle.immutable_csr(ro_src, ro_dst, n, 1 if ntype == "in" else 0)
        res = ref_ffi.graph_uniform_sampling(g, seeds, ntype, hops, fanout, want_graph=True)
        for si, (s, (nodes, edges, layers, adj)) in enumerate(zip(seeds, res)):
            rows = np.repeat(np.arange(len(nodes)), np.diff(adj[0]))
            assert np.array_equal(adj[2], np.arange(len(edges)))
            oracle.check_neighbor_subgraph(*csr, s, fanout, hops, nodes, layers, edges, rows, adj[1])
            if want is not None and fanout >= 10 ** 6:
                w_nodes, w_edges, w_layers, w_adj = want[ci][si]
                assert np.array_equal(nodes, w_nodes) and np.array_equal(edges, w_edges)
                assert np.array_equal(layers, w_layers)
                assert all(np.array_equal(a, b) for a, b in zip(adj, w_adj))
    return True


def main():
    mode = sys.argv[1]
    lib = load(mode)
    # a read-only (immutable) graph of the reference for the sampler: simple graph, hub at 3
    rng_s = np.random.RandomState(4)
    ro_n = 400
    rs, rd = rng_s.randint(0, ro_n, 6000), rng_s.randint(0, ro_n - 30, 6000)
    rd[rng_s.choice(6000, 700, replace=False)] = 3
    key = np.unique(rs * ro_n + rd)
    ro_src, ro_dst = (key // ro_n).astype(np.int64), (key % ro_n).astype(np.int64)
    ro = ref_ffi.RefGraph(ro_src, ro_dst, ro_n, readonly=True, multigraph=False)
    cases = sampling_cases(ro_n, rng_s)
    check_sampling(ro, ro_src, ro_dst, ro_n, cases, None)                  # the unmodified reference passes
    want_samples = [ref_ffi.graph_uniform_sampling(ro, s, t, h, f, want_graph=True) for s, t, f, h in cases]
    ro_vids = np.array([3, 0, 17, 399], np.int64)
    want_ro = queries(ro, ro_n, ro_vids, readonly=True)
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
    assert k >= 31, (k, lib.dglb200_last_error())            # 5 + 12 queries + 5 mutators + 8 sampling + bucketing
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
                     lambda: ref_ffi.graph_find_edges(g, np.array([0, 10 ** 7], np.int64)),
                     lambda: ref_ffi.graph_edges_of_one(g, n + 1),
                     lambda: ref_ffi.graph_degree_of_one(g, -2, out=True),
                     lambda: ref_ffi.graph_edge_ids(g, np.array([0, 1, 2], np.int64), np.array([0, 1], np.int64)),
                     lambda: ref_ffi.graph_edge_ids(g, np.array([0], np.int64), np.array([n + 3], np.int64)),
                     lambda: ref_ffi.get_global("graph_index._CAPI_DGLGraphGetAdj")(g.handle, False, "xyz")):
        try:
            bad_call()
        except ref_ffi.RefError:
            pass
        else:
            raise AssertionError("an invalid query did not raise")
    del g, tiny            # _CAPI_DGLGraphFree (wrapped) drops the device index
    print("PACKED_GRAPH_OK")
    # the neighbour sampler behind the reference's own global names, on the read-only graph; the six index
    # queries pass a read-only handle through to the reference's originals (other slot orders)
    l1 = lib.dglb200_launch_count()
    check_sampling(ro, ro_src, ro_dst, ro_n, cases, want_samples)
    assert lib.dglb200_launch_count() > l1 + 50, "the sampling overrides launched no kernel"
    same(want_ro, queries(ro, ro_n, ro_vids, readonly=True), "read-only graph (pass-through)")
    for bad_call in (lambda: ref_ffi.graph_uniform_sampling(ro, [np.array([ro_n + 1], np.int64)], "in", 1, 2),
                     lambda: ref_ffi.graph_uniform_sampling(ref_ffi.RefGraph(ro_src, ro_dst, ro_n), [ro_vids], "in", 1, 2)):
        try:
            bad_call()
        except ref_ffi.RefError:
            pass
        else:
            raise AssertionError("an invalid sampling call did not raise")
    check_sampling(ro, ro_src, ro_dst, ro_n, cases[:2], want_samples)       # the state survived the failed call
    del ro
    print("PACKED_SAMPLING_OK")
    # degree bucketing: the override against the reference's raw output taken BEFORE the override
    ours = ref_ffi.degree_bucketing(*bucket_in)
    for a, b in zip(oracle.canonicalize_bucketing(*bucket_ref), ours):
        assert np.array_equal(a, b)
    for a, b in zip(ours, oracle.degree_bucketing(*bucket_in)):
        assert np.array_equal(a, b)
    print("BUCKETING_OK")


if __name__ == "__main__":
    main()
