"""-m "not gpu": pins the oracle's neighbour-sampling restatement and its property checker on REAL outputs of
the unmodified reference (tests/golden/sampling_*.npz, made by oracle/gen_golden_sampling.py):
  * "exact" records (fanout covers every degree - the reference's only deterministic case): the restatement
    reproduces them bit for bit;
  * every record, including the reference's random ones: check_neighbor_subgraph accepts it;
  * the checker is not vacuous: each kind of corruption of a real output is rejected.
"""
import glob
import os

import numpy as np
import pytest

from oracle import oracle

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
FILES = sorted(glob.glob(os.path.join(GOLDEN, "sampling_*.npz")), key=lambda p: int(os.path.basename(p)[9:-4]))


def _load(f):
    z = np.load(f)
    n, ntype = int(z["n"]), str(z["ntype"])
    csr = oracle.immutable_csr(z["src"], z["dst"], n, 1 if ntype == "in" else 0)
    rows, cols = (z["sub_dst"], z["sub_src"]) if ntype == "in" else (z["sub_src"], z["sub_dst"])
    return z, csr, rows, cols


def test_all_reference_outputs_pass_the_checker_and_exact_ones_match_the_restatement():
    assert len(FILES) == 32
    exact = 0
    for f in FILES:
        z, csr, rows, cols = _load(f)
        fanout, hops = int(z["fanout"]), int(z["hops"])
        oracle.check_neighbor_subgraph(*csr, z["seeds"], fanout, hops, z["induced_nodes"], z["layer_ids"],
                                       z["induced_edges"], rows, cols)
        rng = None if int(z["exact"]) else np.random.RandomState(4)
        got = oracle.neighbor_sample_subgraph(*csr, z["seeds"], fanout, hops, rng)
        got_rows = np.repeat(np.arange(len(got[0])), np.diff(got[3]))
        oracle.check_neighbor_subgraph(*csr, z["seeds"], fanout, hops, got[0], got[1], got[2], got_rows, got[4])
        if int(z["exact"]):
            assert np.array_equal(got[0], z["induced_nodes"]) and np.array_equal(got[1], z["layer_ids"])
            assert np.array_equal(got[2], z["induced_edges"])
            assert np.array_equal(got_rows, rows) and np.array_equal(got[4], cols)
            exact += 1
    assert exact == 16


def test_checker_rejects_corrupted_outputs():
    f = [p for p in FILES if not int(np.load(p)["exact"]) and int(np.load(p)["hops"]) >= 2][0]
    z, csr, rows, cols = _load(f)
    fanout, hops = int(z["fanout"]), int(z["hops"])
    base = dict(nodes=z["induced_nodes"].copy(), layers=z["layer_ids"].copy(), edges=z["induced_edges"].copy(),
                rows=rows.copy(), cols=cols.copy())

    def run(**kw):
        a = dict(base)
        a.update(kw)
        oracle.check_neighbor_subgraph(*csr, z["seeds"], fanout, hops, a["nodes"], a["layers"], a["edges"],
                                       a["rows"], a["cols"])
    run()
    indptr, indices, eids = csr
    row0 = int(rows[0])
    k = int((rows == row0).sum())
    bad = base["edges"].copy()
    bad[0] = bad[1] if k > 1 else (bad[0] + 1) % len(eids)               # duplicate / foreign slot
    with pytest.raises(AssertionError):
        run(edges=bad)
    if k > 1:
        swapped = base["edges"].copy()
        swapped[[0, 1]] = swapped[[1, 0]]                                 # not ascending
        sc = base["cols"].copy()
        sc[[0, 1]] = sc[[1, 0]]
        with pytest.raises(AssertionError):
            run(edges=swapped, cols=sc)
    with pytest.raises(AssertionError):                                   # a missing edge
        run(edges=base["edges"][1:], rows=base["rows"][1:], cols=base["cols"][1:])
    lay = base["layers"].copy()
    lay[np.argmax(lay)] = 0                                               # a non-seed claimed as seed
    with pytest.raises(AssertionError):
        run(layers=lay)
    lay = base["layers"].copy()
    i = int(np.argmax(lay == 1))
    lay[i] = 2                                                            # not the BFS level
    with pytest.raises(AssertionError):
        run(layers=lay)
    with pytest.raises(AssertionError):
        run(nodes=base["nodes"][::-1].copy())


def test_live_reference_sampler_against_the_restatement_and_the_checker():
    """The compiled reference (oracle/_ref/libdgl.so) called live through its C ABI: multigraph with hubs,
    4 seed arrays per call (its OpenMP path), both neighbour types; deterministic calls equal the
    restatement bit for bit, random ones pass the checker."""
    from oracle import ref_ffi
    if not ref_ffi.available():
        pytest.skip("oracle/_ref/libdgl.so not built")
    rng = np.random.RandomState(12)
    n, e = 1500, 30000
    src, dst = rng.randint(0, n, e), rng.randint(0, n - 100, e)
    dst[rng.choice(e, 3000, replace=False)] = 7                       # hub; the last 100 have no in-edge
    key = np.unique(src * n + dst)                                    # simple graph: slot order is unique
    src, dst = key // n, key % n
    g = ref_ffi.RefGraph(src, dst, n, readonly=True, multigraph=False)
    for ntype in ("in", "out"):
        csr = oracle.immutable_csr(src, dst, n, 1 if ntype == "in" else 0)
        for fanout, hops in ((4, 2), (1, 3), (2900, 1), (10 ** 6, 2)):
            seeds = [rng.randint(0, n, 30) for _ in range(4)]
            seeds[0][0] = 7
            res = ref_ffi.graph_uniform_sampling(g, seeds, ntype, hops, fanout, want_graph=True)
            for s, (nodes, edges, layers, adj) in zip(seeds, res):
                rows = np.repeat(np.arange(len(nodes)), np.diff(adj[0]))
                assert np.array_equal(adj[2], np.arange(len(edges)))
                oracle.check_neighbor_subgraph(*csr, s, fanout, hops, nodes, layers, edges, rows, adj[1])
                if fanout >= 10 ** 6:
                    want = oracle.neighbor_sample_subgraph(*csr, s, fanout, hops)
                    assert np.array_equal(want[0], nodes) and np.array_equal(want[1], layers)
                    assert np.array_equal(want[2], edges) and np.array_equal(want[3], adj[0])
                    assert np.array_equal(want[4], adj[1])
