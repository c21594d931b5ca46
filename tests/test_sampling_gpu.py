"""-m gpu: neighbour sampling on the device (csrc/sampling.cu, dgl_b200/sampling.py, contrib/sampling).

Parity, in three layers:
  1. tests/golden/sampling_*.npz "exact" records - calls of the UNMODIFIED reference whose fanout covers every
     degree (the only deterministic case: it seeds rand_r with time()) - bit for bit: induced nodes, layer ids,
     induced edges, subgraph edges;
  2. the other records' (graph, seeds, fanout, hops), sampled here, must pass
     oracle.check_neighbor_subgraph - the invariants of SampleSubgraph, pinned on the reference's own random
     outputs in tests/test_oracle_sampling.py;
  3. the distribution: every slot of a row equally likely (chi-square), every pair of slots equally likely,
     on all three branches of the kernel (kept set, dropped set, sequential selection).
Plus the reference's own tests/compute/test_sampler.py, restated on this package's NeighborSampler.
The bodies also run on the host emulation of the kernel sources (tests/test_kernels_host_emu.py).
"""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import oracle

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TRIALS = 1200            # draws per distribution test (the host emulation lowers it)


def device():
    return torch.device("cuda")


def _golden_cases():
    return sorted(glob.glob(os.path.join(GOLDEN, "sampling_*.npz")),
                  key=lambda p: int(os.path.basename(p)[9:-4]))


def _parent(z, dev):
    from dgl_b200 import ops
    n, ntype = int(z["n"]), str(z["ntype"])
    src = torch.from_numpy(z["src"]).to(dev)
    dst = torch.from_numpy(z["dst"]).to(dev)
    r, c = (dst, src) if ntype == "in" else (src, dst)
    csr = ops.csr_build(r, c, n, n, order=1)
    ref_csr = oracle.immutable_csr(z["src"], z["dst"], n, 1 if ntype == "in" else 0)
    return csr, ref_csr, ntype


def test_golden_cases_of_the_reference():
    from dgl_b200 import sampling
    dev = device()
    files = _golden_cases()
    assert len(files) == 32
    exact = 0
    for f in files:
        z = np.load(f)
        csr, (indptr, indices, eids), ntype = _parent(z, dev)
        fanout, hops = int(z["fanout"]), int(z["hops"])
        sub = sampling.sample_subgraph(csr, torch.from_numpy(z["seeds"]).to(dev), fanout, hops, seed=17,
                                       neighbor_type=ntype)
        nodes, layers = sub.induced_nodes.cpu().numpy(), sub.layer_ids.cpu().numpy()
        edges = sub.induced_edges.cpu().numpy()
        rows, cols = sub.sub_rows.cpu().numpy(), sub.sub_cols.cpu().numpy()
        assert np.array_equal(sub.csr.eids.cpu().numpy(), np.arange(len(edges)))
        assert np.array_equal(np.repeat(np.arange(len(nodes)), np.diff(sub.csr.indptr.cpu().numpy())), rows)
        oracle.check_neighbor_subgraph(indptr, indices, eids, z["seeds"], fanout, hops, nodes, layers, edges,
                                       rows, cols)
        if int(z["exact"]):
            want_rows, want_cols = (z["sub_dst"], z["sub_src"]) if ntype == "in" else (z["sub_src"], z["sub_dst"])
            assert np.array_equal(nodes, z["induced_nodes"]), f
            assert np.array_equal(layers, z["layer_ids"]), f
            assert np.array_equal(edges, z["induced_edges"]), f
            assert np.array_equal(rows, want_rows) and np.array_equal(cols, want_cols), f
            s, d = sub.edges()
            assert np.array_equal(s.cpu().numpy(), z["sub_src"]) and np.array_equal(d.cpu().numpy(), z["sub_dst"])
            exact += 1
    assert exact == 16


def _one_row_graph(deg, dev, extra_rows=3):
    """Row 1 has `deg` neighbours (ids 10 .. 10+deg), the other rows are empty or short."""
    from dgl_b200 import ops
    n = deg + 10 + extra_rows
    dst = np.concatenate([np.full(deg, 1), np.array([2, 2])]).astype(np.int64)
    src = np.concatenate([np.arange(10, 10 + deg), np.array([5, 6])]).astype(np.int64)
    csr = ops.csr_build(torch.from_numpy(dst).to(dev), torch.from_numpy(src).to(dev), n, n, order=1)
    return csr, n


@pytest.mark.parametrize("deg,fanout,branch", [(40, 5, "kept set"), (40, 33, "dropped set"), (12, 11, "dropped set"),
                                               (700, 300, "selection"), (300, 128, "kept set"),
                                               (300, 172, "dropped set")])
def test_every_slot_and_every_pair_equally_likely(deg, fanout, branch):
    """`trials` independent seeds on one row: slot frequencies against the uniform law (chi-square, generous
    0.9999 quantile so a correct sampler fails once in 10^4 runs - and the seeds are fixed), pair frequencies
    of the first 6 slots, and per call: `fanout` distinct ascending slots."""
    from dgl_b200 import ops
    dev = device()
    csr, n = _one_row_graph(deg, dev)
    trials = TRIALS // 4 if deg > 256 else TRIALS
    rows = torch.tensor([1, 2, 0], dtype=torch.int64, device=dev)
    counts = np.zeros(deg, dtype=np.int64)
    pairs = np.zeros((6, 6), dtype=np.int64)
    start = int(csr.indptr[1])
    for t in range(trials):
        off, slot, col, eid = ops.neighbor_sample(csr, rows, fanout, seed=1000 + t, want=("slot", "col", "eid"))
        off = off.cpu().numpy()
        assert off.tolist() == [0, fanout, fanout + 2, fanout + 2]
        s = slot.cpu().numpy()
        mine = s[:fanout] - start
        assert np.all(np.diff(mine) > 0) and mine.min() >= 0 and mine.max() < deg
        assert np.array_equal(col.cpu().numpy()[:fanout], csr.indices.cpu().numpy()[s[:fanout]])
        assert np.array_equal(eid.cpu().numpy()[:fanout], csr.eids.cpu().numpy()[s[:fanout]])
        assert s[fanout:].tolist() == [int(csr.indptr[2]), int(csr.indptr[2]) + 1]        # short row: whole
        counts[mine] += 1
        head = mine[mine < 6]
        pairs[np.ix_(head, head)] += 1
    p = fanout / deg
    expect = trials * p
    chi2 = float(((counts - expect) ** 2 / (expect * (1 - p))).sum())      # ~ chi-square, deg - 1 d.o.f.
    dof = deg - 1
    limit = dof + 3.72 * np.sqrt(2 * dof) + 10                             # ~ the 0.9999 quantile, normal approx.
    assert chi2 < limit, (branch, chi2, limit)
    p2 = fanout * (fanout - 1) / (deg * (deg - 1))                         # both of two given slots picked
    off_diag = pairs[~np.eye(6, dtype=bool)]
    sigma = np.sqrt(trials * p2 * (1 - p2)) + 1e-9
    assert np.all(np.abs(off_diag - trials * p2) < 5 * sigma + 3), (branch, off_diag, trials * p2)


def test_same_seed_same_sample_and_seeds_differ():
    from dgl_b200 import ops
    dev = device()
    rng = np.random.RandomState(3)
    n, e = 500, 20000
    src, dst = rng.randint(0, n, e), rng.randint(0, n, e)
    csr = ops.csr_build(torch.from_numpy(dst).to(dev), torch.from_numpy(src).to(dev), n, n, order=1)
    rows = torch.from_numpy(rng.permutation(n)[:300].astype(np.int64)).to(dev)
    a = ops.neighbor_sample(csr, rows, 7, seed=5, want=("slot",))[1]
    b = ops.neighbor_sample(csr, rows, 7, seed=5, want=("slot",))[1]
    # a different launch shape (the rows reversed): per-row result unchanged
    c_off, c = ops.neighbor_sample(csr, rows.flip(0), 7, seed=5, want=("slot",))
    d = ops.neighbor_sample(csr, rows, 7, seed=6, want=("slot",))[1]
    assert torch.equal(a, b)
    assert torch.equal(a.reshape(-1, 7), c.reshape(-1, 7).flip(0))
    assert not torch.equal(a, d)
    from dgl_b200 import DGLError
    with pytest.raises(DGLError):
        ops.neighbor_sample(csr, torch.tensor([0, n], dtype=torch.int64, device=dev), 3, seed=0)


HUB_GRAPH = dict(n=3000, e=40000, hub_deg=2500, fanouts=((10, 2), (200, 2), (2400, 1), (1, 3), (0, 2), (5, 0),
                                                         (10 ** 6, 2)))     # the host emulation shrinks it


def test_hubs_multigraph_and_zero_fanout():
    """Parallel edges, hub rows on every branch, empty rows, fanout 0, hops 0: invariants of the reference."""
    from dgl_b200 import ops, sampling
    from tests.util import random_graph
    dev = device()
    n = HUB_GRAPH["n"]
    src, dst = random_graph(n, HUB_GRAPH["e"], 77, hubs=3, hub_deg=HUB_GRAPH["hub_deg"], zero_deg_frac=0.2)
    src = np.concatenate([src, src[:500]])                    # parallel edges
    dst = np.concatenate([dst, dst[:500]])
    csr = ops.csr_build(torch.from_numpy(dst).to(dev), torch.from_numpy(src).to(dev), n, n, order=1)
    indptr, indices, eids = csr.indptr.cpu().numpy(), csr.indices.cpu().numpy().astype(np.int64), \
        csr.eids.cpu().numpy().astype(np.int64)
    rng = np.random.RandomState(1)
    hub = int(np.argmax(np.diff(indptr)))
    for fanout, hops in HUB_GRAPH["fanouts"]:
        seeds = np.concatenate([rng.randint(0, n, 20), [hub]]).astype(np.int64)
        sub = sampling.sample_subgraph(csr, torch.from_numpy(seeds).to(dev), fanout, hops, seed=fanout + hops)
        oracle.check_neighbor_subgraph(indptr, indices, eids, seeds, fanout, hops, sub.induced_nodes.cpu().numpy(),
                                       sub.layer_ids.cpu().numpy(), sub.induced_edges.cpu().numpy(),
                                       sub.sub_rows.cpu().numpy(), sub.sub_cols.cpu().numpy())
        st = csr._sample_state                                 # reused by every call on this CSR: left clean
        assert int((st.layer != -1).sum()) == 0 and int((st.bitmap != 0).sum()) == 0
        if fanout >= 10 ** 6:                                  # deterministic: equals the oracle's restatement
            want = oracle.neighbor_sample_subgraph(indptr, indices, eids, seeds, fanout, hops)
            assert np.array_equal(sub.induced_nodes.cpu().numpy(), want[0])
            assert np.array_equal(sub.layer_ids.cpu().numpy(), want[1])
            assert np.array_equal(sub.induced_edges.cpu().numpy(), want[2])
            assert np.array_equal(sub.csr.indptr.cpu().numpy(), want[3])
            assert np.array_equal(sub.sub_cols.cpu().numpy(), want[4])


def test_bad_seed_is_an_error_and_the_state_recovers():
    from dgl_b200 import DGLError, ops, sampling
    dev = device()
    rng = np.random.RandomState(8)
    n, e = 400, 6000
    src, dst = rng.randint(0, n, e), rng.randint(0, n, e)
    csr = ops.csr_build(torch.from_numpy(dst).to(dev), torch.from_numpy(src).to(dev), n, n, order=1)
    indptr, indices, eids = (t.cpu().numpy().astype(np.int64) for t in (csr.indptr, csr.indices, csr.eids))
    good = np.array([5, 9, 5, 77], dtype=np.int64)
    for bad in ([3, n], [-1, 4]):
        with pytest.raises(DGLError):
            sampling.sample_subgraph(csr, torch.tensor(bad, dtype=torch.int64, device=dev), 4, 2)
        sub = sampling.sample_subgraph(csr, torch.from_numpy(good).to(dev), 4, 2, seed=3)
        oracle.check_neighbor_subgraph(indptr, indices, eids, good, 4, 2, sub.induced_nodes.cpu().numpy(),
                                       sub.layer_ids.cpu().numpy(), sub.induced_edges.cpu().numpy(),
                                       sub.sub_rows.cpu().numpy(), sub.sub_cols.cpu().numpy())
    # the same call twice: the same subgraph (counter-based draws, order-independent discovery)
    again = sampling.sample_subgraph(csr, torch.from_numpy(good).to(dev), 4, 2, seed=3)
    assert torch.equal(again.induced_nodes, sub.induced_nodes) and torch.equal(again.induced_edges, sub.induced_edges)
    assert torch.equal(again.layer_ids, sub.layer_ids) and torch.equal(again.csr.indptr, sub.csr.indptr)
    assert torch.equal(again.csr.indices, sub.csr.indices)
    with pytest.raises(DGLError):
        sampling.sample_subgraph(csr, torch.empty(0, dtype=torch.int64, device=dev), 4, 2)


def test_c_abi_too_small_workspace_reports_needed_bytes_and_wipes_the_state():
    """dglb200_sample_subgraph straight through ctypes: a workspace that cannot hold the picks fails with
    needed_bytes > given and leaves layer / bitmap clean; the retry with that size succeeds."""
    import ctypes
    from dgl_b200 import _lib, ops
    dev = device()
    rng = np.random.RandomState(21)
    n, e = 2000, 60000
    src, dst = rng.randint(0, n, e), rng.randint(0, n, e)
    csr = ops.csr_build(torch.from_numpy(dst).to(dev), torch.from_numpy(src).to(dev), n, n, order=1)
    st = ops.SampleState(n, dev)
    seeds = torch.from_numpy(rng.choice(n, 200, replace=False).astype(np.int64)).to(dev)
    lib = _lib.get_lib()
    out = _lib.SampledSubgraphStruct()

    def call(nbytes):
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        rc = lib.dglb200_sample_subgraph(n, ops._ptr(csr.indptr), ops._ptr(csr.indices), ops._ptr(csr.eids), 200,
                                         ops._ptr(seeds), 20, 2, 11, ops._ptr(st.layer), ops._ptr(st.bitmap),
                                         ops._ptr(st.word_prefix), ops._ptr(ws), nbytes, ctypes.byref(out),
                                         ops._stream())
        return rc, ws
    given = 48 << 10                          # holds the seeds and hop 0's offsets, not its 4000 picks
    rc, _ = call(given)
    assert rc == -1 and out.needed_bytes > given and "too small" in _lib.last_error()
    assert int((st.layer != -1).sum()) == 0 and int((st.bitmap != 0).sum()) == 0
    need = int(out.needed_bytes)
    for _ in range(8):                        # a later hop may ask for more again
        rc, ws = call(need)
        if rc == 0:
            break
        assert out.needed_bytes > need
        need = int(out.needed_bytes)
    assert rc == 0 and out.n_nodes > 200 and out.n_edges > 4000 and out.needed_bytes <= need
    assert int((st.layer != -1).sum()) == 0 and int((st.bitmap != 0).sum()) == 0


# ---- the reference's tests/compute/test_sampler.py on this package's NeighborSampler ----------------------

def _rand_graph(n, dev, seed=0):
    import scipy.sparse as sp
    import dgl_b200
    arr = (sp.random(n, n, density=0.1, format="coo", random_state=seed) != 0).astype(np.int64)
    return dgl_b200.DGLGraph(arr, readonly=True).to(dev)


def _verify_subgraph(g, subg, seed_id):
    src, dst, eid = g.in_edges(seed_id)
    child_id = subg.map_to_subgraph_nid(seed_id)
    child_src, child_dst, child_eid = subg.in_edges(child_id)
    child_src = child_src.cpu().numpy()
    assert len(np.unique(child_src)) == len(child_src)         # test_sampler.py:41-44
    assert np.all(np.sort(child_src) == child_src)
    child_src1 = subg.map_to_subgraph_nid(src).cpu().numpy()
    child_src1 = child_src1[child_src1 >= 0]
    for i in child_src:
        assert i in child_src1


def test_reference_test_1neighbor_sampler_all():
    import dgl_b200
    g = _rand_graph(100, device())
    count = 0
    for subg, aux in dgl_b200.contrib.sampling.NeighborSampler(g, 1, 100, neighbor_type="in", num_workers=4,
                                                               return_seed_id=True):
        seed_ids = aux["seeds"]
        assert len(seed_ids) == 1
        src, dst, eid = g.in_edges(seed_ids)
        self_loop = int((src == dst).sum()) == 1
        assert subg.number_of_nodes() == (len(src) if self_loop else len(src) + 1)
        assert subg.number_of_edges() >= len(src)
        child_ids = subg.map_to_subgraph_nid(seed_ids)
        child_src, child_dst, child_eid = subg.in_edges(child_ids)
        child_src1 = subg.map_to_subgraph_nid(src)
        assert int((child_src1 == child_src).sum()) == len(src)
        assert torch.equal(subg.parent_eid[child_eid], eid)
        count += 1
    assert count == 100


def test_reference_test_1neighbor_and_10neighbor_sampler():
    import dgl_b200
    g = _rand_graph(100, device(), seed=1)
    for prefetch in (False, True):
        for subg, aux in dgl_b200.contrib.sampling.NeighborSampler(g, 1, 5, neighbor_type="in", num_workers=4,
                                                                   return_seed_id=True, prefetch=prefetch):
            seed_ids = aux["seeds"]
            assert len(seed_ids) == 1
            assert subg.number_of_nodes() <= 6
            assert subg.number_of_edges() <= 5
            _verify_subgraph(g, subg, seed_ids)
    seeds = np.unique(np.random.RandomState(0).randint(0, g.number_of_nodes(), size=10))
    for seed_nodes in (None, torch.from_numpy(seeds)):
        batches = 0
        for subg, aux in dgl_b200.contrib.sampling.NeighborSampler(g, 10, 5, neighbor_type="in", num_workers=4,
                                                                   seed_nodes=seed_nodes, return_seed_id=True):
            seed_ids = aux["seeds"]
            assert subg.number_of_nodes() <= 6 * len(seed_ids)
            assert subg.number_of_edges() <= 5 * len(seed_ids)
            for s in seed_ids:
                _verify_subgraph(g, subg, s.reshape(1))
            batches += 1
        assert batches == (10 if seed_nodes is None else 1)


def test_reference_test_10neighbor_sampler_all_and_features():
    import dgl_b200
    import dgl_b200.function as fn
    dev = device()
    g = _rand_graph(100, dev, seed=2)
    g.ndata["h"] = torch.arange(100, dtype=torch.float32, device=dev).reshape(-1, 1) * torch.ones(1, 4, device=dev)
    g.edata["w"] = torch.arange(g.number_of_edges(), dtype=torch.float32, device=dev).reshape(-1, 1)
    for subg, aux in dgl_b200.contrib.sampling.NeighborSampler(g, 10, 100, neighbor_type="in", num_workers=4,
                                                               return_seed_id=True, num_hops=2):
        seed_ids = aux["seeds"]
        src, dst, eid = g.in_edges(seed_ids)
        child_ids = subg.map_to_subgraph_nid(seed_ids)
        child_src, child_dst, child_eid = subg.in_edges(child_ids)
        child_src1 = subg.map_to_subgraph_nid(src)
        assert int((child_src1 == child_src).sum()) == len(src)
        # features follow the induced ids; one round of message passing on the sampled graph equals the
        # parent's on every vertex whose whole neighbourhood was kept (layers < 2 with fanout 100)
        subg.copy_from_parent()
        assert torch.equal(subg.ndata["h"][:, 0], subg.parent_nid.float())
        assert torch.equal(subg.edata["w"][:, 0], subg.parent_eid.float())
        subg.update_all(fn.copy_src("h", "m"), fn.sum("m", "o"))
        g.update_all(fn.copy_src("h", "m"), fn.sum("m", "o"))
        inner = subg.layer_ids < 2
        assert torch.equal(subg.ndata["o"][inner], g.ndata["o"][subg.parent_nid[inner]])
        with pytest.raises(dgl_b200.DGLError):
            subg.add_nodes(1)
    with pytest.raises(NotImplementedError):
        mutable = dgl_b200.DGLGraph()
        mutable.add_nodes(3)
        dgl_b200.contrib.sampling.NeighborSampler(mutable, 1, 2)
