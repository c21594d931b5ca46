"""Generate tests/golden/msgpass_*.npz by running the UNMODIFIED reference: the two-phase
send / recv entry points, push, (u, v)-pair edge specifications, subset apply_nodes / apply_edges
and user initializers (python/dgl/graph.py:1926-2087, 2301-2396, 1800-1924, 1590-1648;
runtime/scheduler.py:28-109, 428-468).

TEST INFRASTRUCTURE ONLY.  Run in the build container (the only place /root/reference exists):

    make -C oracle ref
    DGL_LIBRARY_PATH=oracle/_ref PYTHONPATH=/root/reference/python:. python oracle/gen_golden_msgpass.py

Shapes and scenarios follow the reference's tests/compute/test_basics.py:363-577 (zero-degree
recv / pull, multigraph send, consecutive sends) on the seeded multigraphs of gen_golden.py.
"""
import os
import sys

import numpy as np
import torch

import dgl
import dgl.function as fn

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from gen_golden import F32, OUT, make_graph, random_case, t  # noqa: E402


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


def gen_case(seed):
    n, src, dst, rng = random_case(seed)
    e = len(src)
    d = dict(n=np.int64(n), src=src, dst=dst)
    x = rng.randn(n, 16).astype(F32)
    w = rng.rand(e).astype(F32)
    d["in_x"], d["in_w"] = x, w
    eids_a = rng.choice(e, e // 3, replace=False).astype(np.int64)
    eids_b = rng.choice(e, e // 4, replace=False).astype(np.int64)       # overlaps eids_a
    recv_a = np.concatenate([rng.choice(n - 8, 14, replace=False), [n - 1, n - 4]]).astype(np.int64)
    d["in_eids_a"], d["in_eids_b"], d["in_recv_a"] = eids_a, eids_b, recv_a

    def fresh():
        g = make_graph(src, dst, n)
        g.ndata["h"] = t(x)
        g.edata["w"] = t(w)
        return g

    # 1. send(eids, UDF) + recv(nodes, builtin sum / max): only edges holding a message count
    for red, rf in (("sum", fn.sum), ("max", fn.max)):
        g = fresh()
        g.send(t(eids_a), msg_hw)
        g.recv(t(recv_a), rf("m", "h"))
        d["sendrecv_builtin_%s_h" % red] = g.ndata["h"].numpy().copy()

    # 2. UDF reducer + apply on a recv set with zero-degree nodes and nodes without messages
    g = fresh()
    g.send(t(eids_a), msg_h)
    g.recv(t(recv_a), red_udf, app2)
    d["sendrecv_udf_apply_h"] = g.ndata["h"].numpy().copy()

    # 3. consecutive sends (the later overwrites), partial recv, then recv of everything left
    g = fresh()
    g.send(t(eids_a), msg_h)
    g.send(t(eids_b), msg_3h)
    g.recv(t(recv_a), fn.sum("m", "h"))
    d["two_sends_partial_h"] = g.ndata["h"].numpy().copy()
    d["two_sends_partial_flags"] = g._msg_index.tousertensor().numpy().astype(np.int64)
    g.recv(reduce_func=fn.sum("m", "h"))
    d["two_sends_rest_h"] = g.ndata["h"].numpy().copy()
    assert int(g._msg_index.tousertensor().sum()) == 0

    # 4. send on every edge + recv on every node == update_all
    g = fresh()
    g.send(message_func=msg_hw)
    g.recv(reduce_func=fn.sum("m", "h"))
    d["send_all_recv_all_h"] = g.ndata["h"].numpy().copy()

    # 5. recv with nothing pending downgrades to apply_nodes
    g = fresh()
    g.recv(t(recv_a), fn.sum("m", "h"), app2)
    d["recv_nothing_sent_h"] = g.ndata["h"].numpy().copy()

    # 6. (u, v) pair forms on a multigraph: every parallel edge between the pairs is triggered
    pick = rng.choice(e, 25, replace=False)
    pu, pv = src[pick].copy(), dst[pick].copy()
    d["in_pair_u"], d["in_pair_v"] = pu, pv
    g = fresh()
    g.send_and_recv((t(pu), t(pv)), fn.src_mul_edge("h", "w", "m"), fn.sum("m", "h"))
    d["snr_pairs_h"] = g.ndata["h"].numpy().copy()
    g = fresh()
    g.send((t(pu), t(pv)), msg_h)
    d["send_pairs_flags"] = g._msg_index.tousertensor().numpy().astype(np.int64)
    g.recv(reduce_func=fn.max("m", "h"))
    d["send_pairs_recv_max_h"] = g.ndata["h"].numpy().copy()

    # 7. push
    push_u = np.unique(src)[:9].astype(np.int64)
    d["in_push_u"] = push_u
    for name, mf in (("copy_src", fn.copy_src("h", "m")), ("src_mul_edge", fn.src_mul_edge("h", "w", "m"))):
        g = fresh()
        g.push(t(push_u), mf, fn.sum("m", "h"), app2)
        d["push_%s_apply_h" % name] = g.ndata["h"].numpy().copy()

    # 8. subset apply_nodes / apply_edges
    g = fresh()
    g.apply_nodes(app2, t(recv_a))
    d["apply_nodes_subset_h"] = g.ndata["h"].numpy().copy()
    g.apply_edges(lambda edges: {"w": edges.data["w"] + edges.src["h"][:, 0] - edges.dst["h"][:, 1]}, t(eids_a))
    d["apply_edges_subset_w"] = g.edata["w"].numpy().copy()
    g.apply_edges(lambda edges: {"s": edges.src["h"] + edges.dst["h"]})
    d["apply_edges_all_s"] = g.edata["s"].numpy().copy()

    # 8b. the sparse matrices handed to users (graph.py:2721-2788)
    g = fresh()
    d["adj_dense"] = g.adjacency_matrix().to_dense().numpy().copy()
    d["adj_t_dense"] = g.adjacency_matrix(transpose=True).to_dense().numpy().copy()
    for ty in ("in", "out", "both"):
        d["inc_%s_dense" % ty] = g.incidence_matrix(ty).to_dense().numpy().copy()

    # 9. user initializers: zero-degree pull rows on the bucketing path, and the untouched rows of
    #    a column that a partial write creates
    g = fresh()
    g.set_n_initializer(init2, "h")
    g.pull(t(recv_a), msg_h, red_udf, app2)
    d["pull_init2_h"] = g.ndata["h"].numpy().copy()
    g = fresh()
    g.set_n_initializer(init2, "h1")
    g.send(t(eids_a), msg_h)
    g.recv(t(recv_a), red_udf_new)
    d["recv_newfield_init2_h1"] = g.ndata["h1"].numpy().copy()
    return d


def gen_traversal(seed):
    """Frontier generators (python/dgl/traversal.py -> src/graph/traversal.h) and propagation
    (python/dgl/propagate.py) on a random DAG (forest-like, TreeLSTM-shaped) and on the cyclic
    multigraph of random_case."""
    rng = np.random.RandomState(100 + seed)
    n = 60
    # DAG: every node i > 0 gets 1-3 parents among the nodes below it; edges child -> parent order
    # is arbitrary, parallel edges allowed
    src, dst = [], []
    for i in range(1, n):
        for p in rng.choice(i, size=min(i, rng.randint(1, 4)), replace=True):
            src.append(int(p))
            dst.append(i)
    perm = rng.permutation(len(src))
    relabel = rng.permutation(n)
    src = relabel[np.asarray(src, dtype=np.int64)[perm]]
    dst = relabel[np.asarray(dst, dtype=np.int64)[perm]]
    d = dict(dag_n=np.int64(n), dag_src=src, dag_dst=dst)
    g = make_graph(src, dst, n)

    def pack(tag, frontiers):
        fr = [f.numpy().astype(np.int64) for f in frontiers]
        d[tag + "_ids"] = np.concatenate(fr) if fr else np.zeros(0, np.int64)
        d[tag + "_sections"] = np.asarray([len(f) for f in fr], dtype=np.int64)

    pack("dag_topo", dgl.topological_nodes_generator(g))
    pack("dag_topo_rev", dgl.topological_nodes_generator(g, True))
    roots = np.asarray([int(relabel[0]), int(relabel[1])], dtype=np.int64)
    d["in_dag_bfs_src"] = roots
    pack("dag_bfs", dgl.bfs_nodes_generator(g, t(roots)))
    pack("dag_bfs_edges", dgl.bfs_edges_generator(g, t(roots)))
    x = rng.randn(n, 8).astype(F32)
    d["in_dag_x"] = x
    g.ndata["h"] = t(x)
    dgl.prop_nodes_topo(g, message_func=fn.copy_src("h", "m"), reduce_func=fn.sum("m", "h"))
    d["dag_prop_topo_sum_h"] = g.ndata["h"].numpy().copy()
    g.ndata["h"] = t(x)
    dgl.prop_nodes_topo(g, message_func=msg_h, reduce_func=red_udf, apply_node_func=app2)
    d["dag_prop_topo_udf_h"] = g.ndata["h"].numpy().copy()
    g.ndata["h"] = t(x)
    dgl.prop_edges(g, dgl.bfs_edges_generator(g, t(roots)), fn.copy_src("h", "m"), fn.max("m", "h"))
    d["dag_prop_bfs_edges_max_h"] = g.ndata["h"].numpy().copy()

    # cyclic multigraph: BFS both directions
    cn, csrc, cdst, crng = random_case(seed)
    cg = make_graph(csrc, cdst, cn)
    start = np.asarray([5, 3, 17], dtype=np.int64)
    d["cyc_n"], d["cyc_src"], d["cyc_dst"], d["in_cyc_bfs_src"] = np.int64(cn), csrc, cdst, start
    pack("cyc_bfs", dgl.bfs_nodes_generator(cg, t(start)))
    pack("cyc_bfs_rev", dgl.bfs_nodes_generator(cg, t(start), True))
    pack("cyc_bfs_edges", dgl.bfs_edges_generator(cg, t(start)))
    pack("cyc_bfs_edges_rev", dgl.bfs_edges_generator(cg, t(start), True))
    cx = crng.randn(cn, 8).astype(F32)
    d["in_cyc_x"] = cx
    cg.ndata["h"] = t(cx)
    dgl.prop_nodes_bfs(cg, t(start), message_func=fn.copy_src("h", "m"), reduce_func=fn.sum("m", "h"))
    d["cyc_prop_bfs_sum_h"] = cg.ndata["h"].numpy().copy()
    return d


def gen_batched(seed):
    """dgl.batch + update_all on the union + per-graph readout (python/dgl/batched_graph.py)."""
    rng = np.random.RandomState(200 + seed)
    sizes = [(7, 20), (1, 0), (12, 40), (5, 3), (9, 30)]           # (nodes, edges); one graph has no edge
    d = dict(batch_sizes=np.asarray(sizes, dtype=np.int64))
    graphs = []
    for i, (n, e) in enumerate(sizes):
        src = rng.randint(0, n, size=e).astype(np.int64)
        dst = rng.randint(0, n, size=e).astype(np.int64)
        g = make_graph(src, dst, n)
        x = rng.randn(n, 3, 4).astype(F32)
        w = rng.rand(n).astype(F32)
        ew = rng.randn(e, 5).astype(F32)
        es = rng.rand(e).astype(F32)
        g.ndata["h"], g.ndata["w"] = t(x), t(w)
        if e > 0:
            g.edata["e"], g.edata["s"] = t(ew), t(es)
        graphs.append(g)
        d["g%d_src" % i], d["g%d_dst" % i] = src, dst
        d["in_g%d_h" % i], d["in_g%d_w" % i], d["in_g%d_e" % i], d["in_g%d_s" % i] = x, w, ew, es
    bg = dgl.batch(graphs)
    d["batch_num_nodes"] = np.asarray(bg.batch_num_nodes, dtype=np.int64)
    d["batch_num_edges"] = np.asarray(bg.batch_num_edges, dtype=np.int64)
    bs, bd = bg.edges()
    d["batched_src"], d["batched_dst"] = bs.numpy().astype(np.int64), bd.numpy().astype(np.int64)
    d["batched_h"] = bg.ndata["h"].numpy().copy()
    d["sum_nodes_h"] = dgl.sum_nodes(bg, "h").numpy().copy()
    d["sum_nodes_h_w"] = dgl.sum_nodes(bg, "h", "w").numpy().copy()
    d["mean_nodes_h"] = dgl.mean_nodes(bg, "h").numpy().copy()
    d["mean_nodes_h_w"] = dgl.mean_nodes(bg, "h", "w").numpy().copy()
    d["max_nodes_h"] = dgl.max_nodes(bg, "h").numpy().copy()
    d["sum_edges_e"] = dgl.sum_edges(bg, "e").numpy().copy()
    d["sum_edges_e_s"] = dgl.sum_edges(bg, "e", "s").numpy().copy()
    d["max_edges_e"] = dgl.max_edges(bg, "e").numpy().copy()
    bg.update_all(fn.copy_src("h", "m"), fn.sum("m", "h"))
    d["batched_update_all_h"] = bg.ndata["h"].numpy().copy()
    d["sum_nodes_after_update"] = dgl.sum_nodes(bg, "h").numpy().copy()
    parts = dgl.unbatch(bg)
    for i, gp in enumerate(parts):
        d["unbatch%d_h" % i] = gp.ndata["h"].numpy().copy()
    # gradient of a readout (torch autograd through unsorted_1d_segment_sum in the reference)
    bg2 = dgl.batch(graphs)
    hx = bg2.ndata["h"].clone().requires_grad_(True)
    bg2.ndata["h"] = hx
    gout = rng.randn(len(sizes), 3, 4).astype(F32)
    d["in_readout_gout"] = gout
    (dgl.sum_nodes(bg2, "h", "w") * t(gout)).sum().backward()
    d["grad_sum_nodes_h_w"] = hx.grad.numpy().copy()
    return d


def main():
    os.makedirs(OUT, exist_ok=True)
    for seed in range(2):
        np.savez_compressed(os.path.join(OUT, "batched_%d.npz" % seed), **gen_batched(seed))
    for seed in range(2):
        np.savez_compressed(os.path.join(OUT, "traversal_%d.npz" % seed), **gen_traversal(seed))
    torch.manual_seed(0)
    for seed in range(2):
        np.savez_compressed(os.path.join(OUT, "msgpass_%d.npz" % seed), **gen_case(seed))
    print("wrote", sorted(f for f in os.listdir(OUT) if f.startswith("msgpass")))


if __name__ == "__main__":
    sys.exit(main())
