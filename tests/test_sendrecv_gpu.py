"""-m gpu: the two-phase entry points (send / recv), push, (u, v)-pair edge specifications, subset
apply_nodes / apply_edges and user initializers on the CUDA path, against golden vectors produced
by the unmodified reference (tests/golden/msgpass_*.npz, made by oracle/gen_golden_msgpass.py).
Scenarios follow the reference's tests/compute/test_basics.py:363-577.
fp32 tolerance rtol 1e-5 (scaled atol); message flags bit-exact.

tests/test_host_layer_cpu.py runs the same bodies on CPU tensors with tests/cpu_engine.py behind
`ops` (host logic only)."""
import glob
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = sorted(glob.glob(os.path.join(GOLDEN, "msgpass_*.npz")))
RTOL = 1e-5


def close(got, want, rtol=RTOL):
    want = np.asarray(want)
    got = got.detach().float().cpu().numpy().reshape(want.shape)
    np.testing.assert_allclose(got, want, rtol=rtol, atol=rtol * (np.abs(want).max() + 1e-30))


def cu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def device():
    return torch.device("cuda")


@pytest.fixture(scope="module", params=CASES, ids=[os.path.basename(c) for c in CASES])
def case(request):
    return dict(np.load(request.param))


def fresh(case):
    import dgl_b200
    g = dgl_b200.DGLGraph()
    g.add_nodes(int(case["n"]))
    g.add_edges(torch.from_numpy(case["src"]), torch.from_numpy(case["dst"]))
    g.to(device())
    g.ndata["h"] = cu(case["in_x"])
    g.edata["w"] = cu(case["in_w"])
    return g


def msg_h(edges):
    return {"m": edges.src["h"]}


def msg_hw(edges):
    return {"m": edges.src["h"] * edges.data["w"].unsqueeze(-1)}


def msg_3h(edges):
    return {"m": edges.src["h"] * 3}


def red_udf(nodes):
    return {"h": nodes.data["h"] + nodes.mailbox["m"].sum(1)}


def red_udf_new(nodes):
    return {"h1": nodes.data["h"] + nodes.mailbox["m"].sum(1)}


def app2(nodes):
    return {"h": nodes.data["h"] * 2}


def init2(shape, dtype, ctx, ids):
    return 2 + torch.zeros(shape, dtype=dtype, device=ctx)


def test_send_then_recv_builtin_reducers(case):
    from dgl_b200 import function as fn
    for red, rf in (("sum", fn.sum), ("max", fn.max)):
        g = fresh(case)
        g.send(cu(case["in_eids_a"]), msg_hw)
        g.recv(cu(case["in_recv_a"]), rf("m", "h"))
        close(g.ndata["h"], case["sendrecv_builtin_%s_h" % red])
        names = [ln.split("(")[0].split()[-1] for ln in g.last_prog.lines()]
        assert names[:3] == ["READ_COL", "SPMV", "WRITE_COL_"] and "NODE_UDF" not in names


def test_send_then_recv_udf_reducer_with_apply(case):
    g = fresh(case)
    g.send(cu(case["in_eids_a"]), msg_h)
    g.recv(cu(case["in_recv_a"]), red_udf, app2)
    close(g.ndata["h"], case["sendrecv_udf_apply_h"], rtol=1e-4)


def test_consecutive_sends_partial_recv(case):
    from dgl_b200 import function as fn
    g = fresh(case)
    g.send(cu(case["in_eids_a"]), msg_h)
    g.send(cu(case["in_eids_b"]), msg_3h)            # overwrites the overlap
    g.recv(cu(case["in_recv_a"]), fn.sum("m", "h"))
    close(g.ndata["h"], case["two_sends_partial_h"])
    assert np.array_equal(g.msg_index().cpu().numpy().astype(np.int64), case["two_sends_partial_flags"])
    g.recv(reduce_func=fn.sum("m", "h"))
    close(g.ndata["h"], case["two_sends_rest_h"])
    assert int(g.msg_index().sum()) == 0
    assert not g._msg                          # every message consumed: CLEAR_FRAME_ (scheduler.py:108-109)


def test_send_all_recv_all_equals_update_all(case):
    from dgl_b200 import function as fn
    g = fresh(case)
    g.send(message_func=msg_hw)
    g.recv(reduce_func=fn.sum("m", "h"))
    close(g.ndata["h"], case["send_all_recv_all_h"])
    g2 = fresh(case)
    g2.update_all(fn.src_mul_edge("h", "w", "m"), fn.sum("m", "h"))
    close(g.ndata["h"], g2.ndata["h"].cpu().numpy())
    # builtin message functions are accepted by send as well (the reference's TODO, scheduler.py:42)
    g3 = fresh(case)
    g3.send(message_func=fn.src_mul_edge("h", "w", "m"))
    g3.recv(reduce_func=fn.sum("m", "h"))
    close(g3.ndata["h"], case["send_all_recv_all_h"])


def test_recv_without_pending_messages_is_apply(case):
    from dgl_b200 import function as fn
    g = fresh(case)
    g.recv(cu(case["in_recv_a"]), fn.sum("m", "h"), app2)
    close(g.ndata["h"], case["recv_nothing_sent_h"])


def test_pair_specifications_on_multigraph(case):
    from dgl_b200 import function as fn
    pu, pv = cu(case["in_pair_u"]), cu(case["in_pair_v"])
    g = fresh(case)
    g.send_and_recv((pu, pv), fn.src_mul_edge("h", "w", "m"), fn.sum("m", "h"))
    close(g.ndata["h"], case["snr_pairs_h"])
    g = fresh(case)
    g.send((pu, pv), msg_h)
    assert np.array_equal(g.msg_index().cpu().numpy().astype(np.int64), case["send_pairs_flags"])
    g.recv(reduce_func=fn.max("m", "h"))
    close(g.ndata["h"], case["send_pairs_recv_max_h"])


def test_push_with_apply(case):
    from dgl_b200 import function as fn
    for name, mf in (("copy_src", fn.copy_src("h", "m")), ("src_mul_edge", fn.src_mul_edge("h", "w", "m"))):
        g = fresh(case)
        g.push(cu(case["in_push_u"]), mf, fn.sum("m", "h"), app2)
        close(g.ndata["h"], case["push_%s_apply_h" % name])


def test_subset_apply_nodes_and_edges(case):
    g = fresh(case)
    g.apply_nodes(app2, cu(case["in_recv_a"]))
    close(g.ndata["h"], case["apply_nodes_subset_h"])
    g.apply_edges(lambda edges: {"w": edges.data["w"] + edges.src["h"][:, 0] - edges.dst["h"][:, 1]},
                  cu(case["in_eids_a"]))
    close(g.edata["w"], case["apply_edges_subset_w"])
    g.apply_edges(lambda edges: {"s": edges.src["h"] + edges.dst["h"]})
    close(g.edata["s"], case["apply_edges_all_s"])


def test_user_facing_sparse_matrices(case):
    """DGLGraph.adjacency_matrix / incidence_matrix (graph.py:2721-2788) as dense, against the reference."""
    g = fresh(case)
    assert np.array_equal(g.adjacency_matrix().to_dense().cpu().numpy(), case["adj_dense"])
    assert np.array_equal(g.adjacency_matrix(transpose=True).to_dense().cpu().numpy(), case["adj_t_dense"])
    for ty in ("in", "out", "both"):
        assert np.array_equal(g.incidence_matrix(ty).to_dense().cpu().numpy(), case["inc_%s_dense" % ty]), ty
    g.drop_coo()                                   # answered from the CSRs alone
    assert np.array_equal(g.adjacency_matrix().to_dense().cpu().numpy(), case["adj_dense"])


def test_structure_changes_after_drop_coo(case):
    """add_edges / to() after the edge list was dropped recover it from the CSRs first."""
    from dgl_b200 import function as fn
    g = fresh(case)
    g.in_csr(), g.out_csr()
    g.drop_coo()
    n = g.number_of_nodes()
    g.add_edges(torch.tensor([0, 1]), torch.tensor([2, 2]))
    s, d = g.edges()
    assert np.array_equal(s.cpu().numpy(), np.concatenate([case["src"], [0, 1]]))
    assert np.array_equal(d.cpu().numpy(), np.concatenate([case["dst"], [2, 2]]))
    g.in_csr()
    g.drop_coo()
    g.to(device())
    assert g.number_of_edges() == len(case["src"]) + 2
    g.ndata["h"] = cu(case["in_x"])
    g.update_all(fn.copy_src("h", "m"), fn.sum("m", "y"))
    want = np.zeros_like(case["in_x"])
    np.add.at(want, np.concatenate([case["dst"], [2, 2]]), case["in_x"][np.concatenate([case["src"], [0, 1]])])
    close(g.ndata["y"], want)
    assert n == g.number_of_nodes()


def test_user_initializers(case):
    """tests/compute/test_basics.py:363-513: zero-degree rows on the bucketing path and the untouched
    rows of a column created by a partial write come from the user's initializer."""
    g = fresh(case)
    g.set_n_initializer(init2, "h")
    g.pull(cu(case["in_recv_a"]), msg_h, red_udf, app2)
    close(g.ndata["h"], case["pull_init2_h"], rtol=1e-4)
    g = fresh(case)
    g.set_n_initializer(init2, "h1")
    g.send(cu(case["in_eids_a"]), msg_h)
    g.recv(cu(case["in_recv_a"]), red_udf_new)
    close(g.ndata["h1"], case["recv_newfield_init2_h1"], rtol=1e-4)


def test_recv_0deg_like_reference():
    """tests/compute/test_basics.py:363-398, literally."""
    import dgl_b200
    g = dgl_b200.DGLGraph()
    g.add_nodes(2)
    g.add_edge(0, 1)
    g.to(device())
    g.register_message_func(msg_h)
    g.register_reduce_func(red_udf)
    g.register_apply_node_func(app2)
    g.set_n_initializer(init2, "h")
    old = torch.randn(2, 5, device=device())
    g.ndata["h"] = old
    g.send((0, 1))
    g.recv([0, 1])
    new = g.ndata.pop("h")
    assert torch.allclose(new[0], torch.full((5,), 4.0, device=device()))
    assert torch.allclose(new[1], old.sum(0) * 2)
    old = torch.randn(2, 5, device=device())
    g.ndata["h"] = old
    g.send((0, 1))
    g.recv(0)
    new = g.ndata.pop("h")
    assert torch.allclose(new[0], 2 * old[0])
    assert torch.allclose(new[1], old[1])


def test_send_multigraph_like_reference():
    """tests/compute/test_basics.py:515-577."""
    import dgl_b200
    dev = device()
    g = dgl_b200.DGLGraph(multigraph=True)
    g.add_nodes(3)
    for u, v in ((0, 1), (0, 1), (0, 1), (2, 1)):
        g.add_edge(u, v)
    g.to(dev)

    def message_a(edges):
        return {"a": edges.data["a"]}

    def message_b(edges):
        return {"a": edges.data["a"] * 3}

    def reduce(nodes):
        return {"a": nodes.mailbox["a"].max(1)[0]}

    def answer(*args):
        return torch.stack(args, 0).max(0)[0]

    old = torch.randn(4, 5, device=dev)

    def reset():
        g.ndata["a"] = torch.zeros(3, 5, device=dev)
        g.edata["a"] = old
    reset()
    g.send([0, 2], message_func=message_a)
    g.recv(1, reduce)
    assert torch.allclose(g.ndata["a"][1], answer(old[0], old[2]))
    reset()
    g.send([0, 2, 3], message_func=message_a)
    g.recv(1, reduce)
    assert torch.allclose(g.ndata["a"][1], answer(old[0], old[2], old[3]))
    reset()
    g.send(([0, 2], [1, 1]), message_a)               # every parallel edge 0 -> 1, and 2 -> 1
    g.recv(1, reduce)
    assert torch.allclose(g.ndata["a"][1], old.max(0)[0])
    reset()
    g.send((2, 1), message_a)
    g.send([0, 1], message_func=message_b)
    g.recv(1, reduce)
    assert torch.allclose(g.ndata["a"][1], answer(old[0] * 3, old[1] * 3, old[3]))
    reset()
    g.send(0, message_func=message_a)
    g.send(1, message_func=message_b)
    g.recv(1, reduce)
    assert torch.allclose(g.ndata["a"][1], answer(old[0], old[1] * 3))
    reset()
    g.send_and_recv([0, 2, 3], message_func=message_a, reduce_func=reduce)
    new = g.ndata["a"]
    assert torch.allclose(new[1], answer(old[0], old[2], old[3]))
    assert torch.allclose(new[[0, 2]], torch.zeros(2, 5, device=dev))


# ---- frontier generators + propagation (python/dgl/traversal.py, propagate.py) ------------------

TRAVERSAL = sorted(glob.glob(os.path.join(GOLDEN, "traversal_*.npz")))


@pytest.fixture(scope="module", params=TRAVERSAL, ids=[os.path.basename(c) for c in TRAVERSAL])
def trav(request):
    return dict(np.load(request.param))


def _graph_of(case, tag):
    import dgl_b200
    g = dgl_b200.DGLGraph()
    g.add_nodes(int(case[tag + "_n"]))
    g.add_edges(torch.from_numpy(case[tag + "_src"]), torch.from_numpy(case[tag + "_dst"]))
    return g.to(device())


def _same_frontiers(frontiers, case, tag):
    """Bit-exact: same number of frontiers, same ids in the same order inside each."""
    got = [f.cpu().numpy() for f in frontiers]
    assert [len(f) for f in got] == case[tag + "_sections"].tolist(), tag
    cat = np.concatenate(got) if got else np.zeros(0, np.int64)
    assert np.array_equal(cat, case[tag + "_ids"]), tag


def test_frontier_generators_bit_exact(trav):
    import dgl_b200
    g = _graph_of(trav, "dag")
    _same_frontiers(dgl_b200.topological_nodes_generator(g), trav, "dag_topo")
    _same_frontiers(dgl_b200.topological_nodes_generator(g, reverse=True), trav, "dag_topo_rev")   # the reference's keyword (traversal.py:91)
    roots = cu(trav["in_dag_bfs_src"])
    _same_frontiers(dgl_b200.bfs_nodes_generator(g, roots), trav, "dag_bfs")
    _same_frontiers(dgl_b200.bfs_edges_generator(g, roots), trav, "dag_bfs_edges")
    c = _graph_of(trav, "cyc")
    start = cu(trav["in_cyc_bfs_src"])
    _same_frontiers(dgl_b200.bfs_nodes_generator(c, start), trav, "cyc_bfs")
    _same_frontiers(dgl_b200.bfs_nodes_generator(c, start, reverse=True), trav, "cyc_bfs_rev")
    _same_frontiers(dgl_b200.bfs_edges_generator(c, start), trav, "cyc_bfs_edges")
    _same_frontiers(dgl_b200.bfs_edges_generator(c, start, True), trav, "cyc_bfs_edges_rev")
    from dgl_b200.base import DGLError
    with pytest.raises(DGLError):
        dgl_b200.topological_nodes_generator(c)       # loop detected (traversal.h:209-211)


def test_propagation_against_reference(trav):
    import dgl_b200
    from dgl_b200 import function as fn
    g = _graph_of(trav, "dag")
    g.ndata["h"] = cu(trav["in_dag_x"])
    dgl_b200.prop_nodes_topo(g, message_func=fn.copy_src("h", "m"), reduce_func=fn.sum("m", "h"))
    close(g.ndata["h"], trav["dag_prop_topo_sum_h"], rtol=1e-4)
    g.ndata["h"] = cu(trav["in_dag_x"])
    dgl_b200.prop_nodes_topo(g, message_func=msg_h, reduce_func=red_udf, apply_node_func=app2)
    close(g.ndata["h"], trav["dag_prop_topo_udf_h"], rtol=1e-4)
    g.ndata["h"] = cu(trav["in_dag_x"])
    dgl_b200.prop_edges(g, dgl_b200.bfs_edges_generator(g, cu(trav["in_dag_bfs_src"])),
                        fn.copy_src("h", "m"), fn.max("m", "h"))
    close(g.ndata["h"], trav["dag_prop_bfs_edges_max_h"])
    c = _graph_of(trav, "cyc")
    c.ndata["h"] = cu(trav["in_cyc_x"])
    dgl_b200.prop_nodes_bfs(c, cu(trav["in_cyc_bfs_src"]), message_func=fn.copy_src("h", "m"),
                            reduce_func=fn.sum("m", "h"))
    close(c.ndata["h"], trav["cyc_prop_bfs_sum_h"], rtol=1e-4)


# ---- batched graphs + per-graph readout (python/dgl/batched_graph.py) -------------------------------

BATCHED = sorted(glob.glob(os.path.join(GOLDEN, "batched_*.npz")))


@pytest.fixture(scope="module", params=BATCHED, ids=[os.path.basename(c) for c in BATCHED])
def bcase(request):
    return dict(np.load(request.param))


def _component_graphs(case):
    import dgl_b200
    graphs = []
    for i, (n, e) in enumerate(case["batch_sizes"].tolist()):
        g = dgl_b200.DGLGraph()
        g.add_nodes(n)
        if e > 0:
            g.add_edges(torch.from_numpy(case["g%d_src" % i]), torch.from_numpy(case["g%d_dst" % i]))
        g.to(device())
        g.ndata["h"], g.ndata["w"] = cu(case["in_g%d_h" % i]), cu(case["in_g%d_w" % i])
        if e > 0:
            g.edata["e"], g.edata["s"] = cu(case["in_g%d_e" % i]), cu(case["in_g%d_s" % i])
        graphs.append(g)
    return graphs


def test_batch_readout_unbatch_against_reference(bcase):
    import dgl_b200
    from dgl_b200 import batched_graph as bgm
    from dgl_b200 import function as fn
    from dgl_b200.base import DGLError
    bg = bgm.batch(_component_graphs(bcase))
    assert bg.batch_size == len(bcase["batch_sizes"])
    assert bg.batch_num_nodes == bcase["batch_num_nodes"].tolist()
    assert bg.batch_num_edges == bcase["batch_num_edges"].tolist()
    s, d = bg.edges()
    assert np.array_equal(s.cpu().numpy(), bcase["batched_src"]) and np.array_equal(d.cpu().numpy(), bcase["batched_dst"])
    close(bg.ndata["h"], bcase["batched_h"])
    with pytest.raises(DGLError):
        bg.add_nodes(1)
    close(bgm.sum_nodes(bg, "h"), bcase["sum_nodes_h"])
    close(bgm.sum_nodes(bg, "h", "w"), bcase["sum_nodes_h_w"])
    close(bgm.mean_nodes(bg, "h"), bcase["mean_nodes_h"])
    close(bgm.mean_nodes(bg, "h", "w"), bcase["mean_nodes_h_w"])
    assert np.array_equal(bgm.max_nodes(bg, "h").cpu().numpy(), bcase["max_nodes_h"])     # max is bit-exact
    close(bgm.sum_edges(bg, "e"), bcase["sum_edges_e"])
    close(bgm.sum_edges(bg, "e", "s"), bcase["sum_edges_e_s"])
    assert np.array_equal(bgm.max_edges(bg, "e").cpu().numpy(), bcase["max_edges_e"])
    bg.update_all(fn.copy_src("h", "m"), fn.sum("m", "h"))
    close(bg.ndata["h"], bcase["batched_update_all_h"])
    close(bgm.sum_nodes(bg, "h"), bcase["sum_nodes_after_update"])
    parts = bgm.unbatch(bg)
    assert [p.number_of_nodes() for p in parts] == bcase["batch_num_nodes"].tolist()
    assert [p.number_of_edges() for p in parts] == bcase["batch_num_edges"].tolist()
    for i, p in enumerate(parts):
        close(p.ndata["h"], bcase["unbatch%d_h" % i])
        if p.number_of_edges():
            ps, pd = p.edges()
            assert np.array_equal(ps.cpu().numpy(), bcase["g%d_src" % i]) and np.array_equal(pd.cpu().numpy(), bcase["g%d_dst" % i])
    # a single (unbatched) graph reads out over all of its nodes (batched_graph.py:397-399)
    g0 = _component_graphs(bcase)[0]
    close(bgm.sum_nodes(g0, "h"), bcase["sum_nodes_h"][0])
    close(bgm.max_nodes(g0, "h"), bcase["max_nodes_h"][0])


def test_readout_gradient_against_reference(bcase):
    from dgl_b200 import batched_graph as bgm
    bg = bgm.batch(_component_graphs(bcase))
    hx = bg.ndata["h"].clone().requires_grad_(True)
    bg.ndata["h"] = hx
    (bgm.sum_nodes(bg, "h", "w") * cu(bcase["in_readout_gout"])).sum().backward()
    close(hx.grad, bcase["grad_sum_nodes_h_w"])
    # max readout routes the gradient to the arg-max node of every (graph, feature)
    hx2 = bg.ndata["h"].detach().clone().requires_grad_(True)
    bg.ndata["h"] = hx2
    bgm.max_nodes(bg, "h").sum().backward()
    ref = bg.ndata["h"].detach().clone().requires_grad_(True)
    off, tot = 0, 0
    for n_i in bg.batch_num_nodes:
        if n_i:
            tot = tot + ref[off:off + n_i].max(0)[0].sum()
        off += n_i
    tot.backward()
    close(hx2.grad, ref.grad.cpu().numpy())
