"""Generate tests/golden/*.npz + plans.json by running the UNMODIFIED reference.

TEST INFRASTRUCTURE ONLY.  Run in the build container (the only place /root/reference exists):

    make -C oracle ref
    DGL_LIBRARY_PATH=oracle/_ref PYTHONPATH=/root/reference/python python oracle/gen_golden.py

It imports the reference's Python package (`dgl` 0.1.3, PyTorch backend, CPU) on top of the
reference C++ compiled by oracle/Makefile into oracle/_ref/libdgl.so, drives the reference's own
public API (DGLGraph.update_all / send_and_recv / pull, dgl.function, the GAT example's edge path,
_CAPI_DGLGraphGetAdj, _CAPI_DGLDegreeBucketing, Prog.pprint) on seeded toy inputs, and stores
inputs + outputs.  The committed vectors pin oracle/oracle.py, oracle/oracle.c and - through the
`-m gpu` tests - the CUDA path.  Shapes follow the reference's tests
(tests/compute/test_specialization.py:9-21,361-370,517-569; test_function.py:5-49).
"""
import contextlib
import io
import json
import os
import sys

import numpy as np
import torch

import dgl
import dgl.function as fn
from dgl import utils
from dgl.graph_index import _CAPI_DGLGraphGetAdj
from dgl.runtime import ir, scheduler
from dgl.runtime.degree_bucketing import _CAPI_DGLDegreeBucketing

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")


def make_graph(src, dst, n, readonly=False):
    if readonly:
        g = dgl.DGLGraph(list(zip(src.tolist(), dst.tolist())), readonly=True)
        assert g.number_of_nodes() == n
        return g
    g = dgl.DGLGraph()
    g.add_nodes(n)
    g.add_edges(torch.from_numpy(src), torch.from_numpy(dst))
    return g


def random_case(seed):
    """Multigraph with parallel edges, self loops, a hub and zero-in-degree nodes."""
    rng = np.random.RandomState(seed)
    n = 40 + 10 * seed
    e = 300 + 50 * seed
    src = rng.randint(0, n, size=e).astype(np.int64)
    dst = rng.randint(0, n - 8, size=e).astype(np.int64)       # last 8 nodes: no in-edges
    dst[rng.choice(e, 60, replace=False)] = 3                   # hub
    src[-1], dst[-1] = n - 1, 0                                 # make node n-1 exist in readonly graphs
    return n, src, dst, rng


def t(a):
    return torch.from_numpy(np.ascontiguousarray(a))


def run_update_all(g, mfunc, rfunc, out_field):
    g.update_all(mfunc, rfunc)
    return g.ndata[out_field].detach().numpy().copy()


def adjacency(g, transpose):
    rst = _CAPI_DGLGraphGetAdj(g._graph._handle, transpose, "csr")
    return [rst(i).asnumpy().astype(np.int64) for i in range(3)]


def bucketing(mids, vids, recv):
    b = _CAPI_DGLDegreeBucketing(utils.toindex(t(mids)).todgltensor(),
                                 utils.toindex(t(vids)).todgltensor(),
                                 utils.toindex(t(recv)).todgltensor())
    return [b(i).asnumpy().astype(np.int64) for i in range(5)]


def plan_text(build):
    buf = io.StringIO()
    with ir.prog() as prog:
        build()
        with contextlib.redirect_stdout(buf):
            prog.pprint()
    return buf.getvalue().strip().splitlines()


F32 = np.float32


def gen_case(seed):
    n, src, dst, rng = random_case(seed)
    e = len(src)
    d = dict(n=np.int64(n), src=src, dst=dst)
    g = make_graph(src, dst, n)

    # ---- structure: Graph::GetAdj csr (both orientations), ImmutableGraph CSR, degrees ----------
    for tr in (False, True):
        ip, ix, ei = adjacency(g, tr)
        tag = "adj_csr_t" if tr else "adj_csr"
        d[tag + "_indptr"], d[tag + "_indices"], d[tag + "_eid"] = ip, ix, ei
    gi = make_graph(src, dst, n, readonly=True)
    ip, ix, ei = adjacency(gi, False)
    d["imm_csr_indptr"], d["imm_csr_indices"], d["imm_csr_eid"] = ip, ix, ei
    d["in_degrees"] = g.in_degrees().numpy().astype(np.int64)
    d["out_degrees"] = g.out_degrees().numpy().astype(np.int64)

    # ---- update_all, every builtin pair, 1-D / 2-D / 3-D features ------------------------------
    feats = {"f1": rng.randn(n).astype(F32), "f3": rng.randn(n, 3).astype(F32),
             "f55": rng.randn(n, 5, 5).astype(F32), "f64": rng.randn(n, 64).astype(F32),
             "f46": rng.randn(n, 4, 6).astype(F32)}
    wts = {"ws": rng.rand(e).astype(F32), "w1": rng.rand(e, 1).astype(F32),
           "w55": rng.rand(e, 5, 5).astype(F32), "w41": rng.rand(e, 4, 1).astype(F32),
           "w64": rng.randn(e, 64).astype(F32)}
    for k, v in list(feats.items()) + list(wts.items()):
        d["in_" + k] = v
    for fk, x in feats.items():
        for red, rf in (("sum", fn.sum), ("max", fn.max)):
            g.ndata["h"] = t(x)
            d["copy_src_%s_%s" % (red, fk)] = run_update_all(g, fn.copy_src("h", "m"), rf("m", "y"), "y")
    for fk, wk in (("f1", "ws"), ("f3", "ws"), ("f3", "w1"), ("f64", "ws"), ("f64", "w1"),
                   ("f55", "ws"), ("f55", "w55"), ("f46", "w41"), ("f64", "w64")):
        for red, rf in (("sum", fn.sum), ("max", fn.max)):
            g.ndata["h"] = t(feats[fk])
            g.edata["w"] = t(wts[wk])
            d["src_mul_edge_%s_%s_%s" % (red, fk, wk)] = run_update_all(
                g, fn.src_mul_edge("h", "w", "m"), rf("m", "y"), "y")
    for wk in ("ws", "w1", "w55", "w64"):
        for red, rf in (("sum", fn.sum), ("max", fn.max)):
            g.edata["w"] = t(wts[wk])
            d["copy_edge_%s_%s" % (red, wk)] = run_update_all(g, fn.copy_edge("w", "m"), rf("m", "y"), "y")

    # ---- gradients through update_all (torch autograd of sparse.mm / of the UDF fallbacks) -----
    dy = rng.randn(n, 64).astype(F32)
    d["in_dy64"] = dy
    x = t(feats["f64"]).requires_grad_(True)
    g.ndata["h"] = x
    g.update_all(fn.copy_src("h", "m"), fn.sum("m", "y"))
    g.ndata["y"].backward(t(dy))
    d["grad_copy_src_sum_f64_dx"] = x.grad.numpy().copy()
    for wk in ("ws", "w64"):
        x = t(feats["f64"]).requires_grad_(True)
        w = t(wts[wk]).requires_grad_(True)
        g.ndata["h"] = x
        g.edata["w"] = w
        g.update_all(fn.src_mul_edge("h", "w", "m"), fn.sum("m", "y"))
        g.ndata["y"].backward(t(dy))
        d["grad_src_mul_edge_sum_f64_%s_dx" % wk] = x.grad.numpy().copy()
        d["grad_src_mul_edge_sum_f64_%s_dw" % wk] = w.grad.numpy().copy()
    x = t(feats["f64"]).requires_grad_(True)
    g.ndata["h"] = x
    g.update_all(fn.copy_src("h", "m"), fn.max("m", "y"))
    g.ndata["y"].backward(t(dy))
    d["grad_copy_src_max_f64_dx"] = x.grad.numpy().copy()

    # ---- edge / node subsets: send_and_recv, pull (rows in unique-ascending dst order) ---------
    eids = rng.choice(e, e // 3, replace=False).astype(np.int64)
    pull_nodes = np.concatenate([rng.choice(n - 8, 12, replace=False), [n - 1, n - 2]]).astype(np.int64)
    d["in_snr_eids"], d["in_pull_nodes"] = eids, pull_nodes
    for name, mf in (("copy_src", fn.copy_src("h", "m")), ("src_mul_edge", fn.src_mul_edge("h", "w", "m"))):
        for red, rf in (("sum", fn.sum), ("max", fn.max)):
            g.ndata["h"] = t(feats["f64"])
            g.edata["w"] = t(wts["ws"])
            g.send_and_recv(t(eids), mf, rf("m", "h"))          # overwrite an existing column
            d["snr_%s_%s_h" % (name, red)] = g.ndata["h"].numpy().copy()
            g.ndata["h"] = t(feats["f64"])
            g.pull(t(pull_nodes), mf, rf("m", "h"))
            d["pull_%s_%s_h" % (name, red)] = g.ndata["h"].numpy().copy()

    # ---- GAT example edge path (examples/pytorch/gat/train.py:61-97, dropout off) --------------
    heads, hid = 4, 6
    ft = rng.randn(n, heads, hid).astype(F32)
    a1 = rng.randn(n, heads, 1).astype(F32)
    a2 = rng.randn(n, heads, 1).astype(F32)
    d["in_gat_ft"], d["in_gat_a1"], d["in_gat_a2"] = ft, a1, a2
    gg = make_graph(np.concatenate([src, np.arange(n)]), np.concatenate([dst, np.arange(n)]), n)
    gg.ndata.update({"ft": t(ft), "a1": t(a1), "a2": t(a2)})

    def edge_attention(edges):
        a = torch.nn.functional.leaky_relu(edges.src["a1"] + edges.dst["a2"], 0.2)
        a = torch.exp(a).clamp(-10, 10)
        return {"a": a, "a_drop": a}
    gg.apply_edges(edge_attention)
    gg.update_all([fn.src_mul_edge("ft", "a_drop", "ft"), fn.copy_edge("a", "a")],
                  [fn.sum("ft", "ft"), fn.sum("a", "z")])
    d["gat_a"] = gg.edata["a"].numpy().copy()
    d["gat_z"] = gg.ndata["z"].numpy().copy()
    d["gat_out"] = (gg.ndata["ft"] / gg.ndata["z"]).numpy().copy()

    # ---- degree bucketing, raw reference order (canonicalised by the tests) --------------------
    mids = np.arange(e, dtype=np.int64)
    recv = np.unique(np.concatenate([dst, [n - 1, n - 3]])).astype(np.int64)
    for i, arr in enumerate(bucketing(mids, dst, recv)):
        d["bucket_full_%d" % i] = arr
    d["in_bucket_full_recv"] = recv
    sub = rng.choice(e, e // 4, replace=False).astype(np.int64)
    sub_recv = np.unique(dst[sub]).astype(np.int64)
    for i, arr in enumerate(bucketing(sub, dst[sub], sub_recv)):
        d["bucket_sub_%d" % i] = arr
    d["in_bucket_sub_mids"], d["in_bucket_sub_recv"] = sub, sub_recv
    return d


def gen_hub():
    """The 10-node / 17-edge graph and literal answers of tests/compute/test_function.py:5-49,
    re-derived by running the reference (the test's literals are asserted here)."""
    g = dgl.DGLGraph()
    g.add_nodes(10)
    for i in range(1, 9):
        g.add_edge(0, i)
        g.add_edge(i, 9)
    g.add_edge(9, 0)
    h = torch.tensor([1., 2., 3., 4., 5., 6., 7., 8., 9., 10.])
    eh = torch.tensor([1., 2., 1., 3., 1., 4., 1., 5., 1., 6., 1., 7., 1., 8., 1., 9., 10.])
    src, dst = g.edges()
    d = dict(src=src.numpy().astype(np.int64), dst=dst.numpy().astype(np.int64), h=h.numpy(), eh=eh.numpy())
    g.ndata["h"] = h
    g.edata["h"] = eh
    g.update_all(fn.copy_src("h", "m"), fn.sum("m", "o"))
    d["copy_src"] = g.ndata["o"].numpy().copy()
    assert d["copy_src"].tolist() == [10., 1., 1., 1., 1., 1., 1., 1., 1., 44.]
    g.update_all(fn.copy_edge("h", "m"), fn.sum("m", "o"))
    d["copy_edge"] = g.ndata["o"].numpy().copy()
    assert d["copy_edge"].tolist() == [10., 1., 1., 1., 1., 1., 1., 1., 1., 44.]
    g.update_all(fn.src_mul_edge("h", "h", "m"), fn.sum("m", "o"))
    d["src_mul_edge"] = g.ndata["o"].numpy().copy()
    assert d["src_mul_edge"].tolist() == [100., 1., 1., 1., 1., 1., 1., 1., 1., 284.]
    return d


def gen_plans():
    """Prog.pprint() text of the reference for calls whose lowering the engine keeps identical."""
    g = dgl.DGLGraph()
    g.add_nodes(4)
    g.add_edges([0, 1, 2, 0], [1, 2, 3, 3])
    g.ndata["h"] = torch.zeros(4, 2)
    g.edata["w"] = torch.zeros(4)
    apply_fn = lambda nodes: {"h": nodes.data["h"] * 2}   # noqa: E731
    plans = {}
    plans["update_all_copy_src_sum"] = plan_text(lambda: scheduler.schedule_update_all(
        g, fn.copy_src("h", "m"), fn.sum("m", "h"), None))
    plans["update_all_src_mul_edge_sum"] = plan_text(lambda: scheduler.schedule_update_all(
        g, fn.src_mul_edge("h", "w", "m"), fn.sum("m", "h"), None))
    plans["update_all_copy_src_sum_apply"] = plan_text(lambda: scheduler.schedule_update_all(
        g, fn.copy_src("h", "m"), fn.sum("m", "h"), apply_fn))
    plans["update_all_two_pairs"] = plan_text(lambda: scheduler.schedule_update_all(
        g, [fn.copy_src("h", "m1"), fn.src_mul_edge("h", "w", "m2")],
        [fn.sum("m1", "h1"), fn.sum("m2", "h2")], None))
    plans["update_all_udf_msg_sum"] = plan_text(lambda: scheduler.schedule_update_all(
        g, lambda edges: {"m": edges.src["h"]}, fn.sum("m", "h"), None))
    u, v, eid = g._graph.find_edges(utils.toindex(torch.tensor([0, 2])))
    plans["snr_copy_src_sum"] = plan_text(lambda: scheduler.schedule_snr(
        g, (u, v, eid), fn.copy_src("h", "m"), fn.sum("m", "h"), None, False))
    plans["pull_copy_src_sum"] = plan_text(lambda: scheduler.schedule_pull(
        g, utils.toindex(torch.tensor([3, 1])), fn.copy_src("h", "m"), fn.sum("m", "h"), None, False))
    plans["apply_nodes"] = plan_text(lambda: scheduler.schedule_apply_nodes(
        g, utils.toindex(slice(0, 4)), apply_fn, False))
    return plans


def main():
    os.makedirs(OUT, exist_ok=True)
    torch.manual_seed(0)
    for seed in range(3):
        np.savez_compressed(os.path.join(OUT, "case_%d.npz" % seed), **gen_case(seed))
    np.savez_compressed(os.path.join(OUT, "hub.npz"), **gen_hub())
    with open(os.path.join(OUT, "plans.json"), "w") as f:
        json.dump(gen_plans(), f, indent=1, sort_keys=True)
    meta = {"reference": "GaiYu0/dgl-1 (dgl %s)" % dgl.__version__, "torch": torch.__version__,
            "numpy": np.__version__, "generator": "oracle/gen_golden.py", "seeds": list(range(3))}
    with open(os.path.join(OUT, "META.json"), "w") as f:
        json.dump(meta, f, indent=1, sort_keys=True)
    print("wrote", sorted(os.listdir(OUT)))


if __name__ == "__main__":
    sys.exit(main())
