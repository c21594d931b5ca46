"""Per-config kernel timings for the BASELINE.json configs that are not bench.py's line
(C1 Cora-shaped GCN, C3 PubMed-shaped GAT, C4 AIFB-shaped R-GCN) plus the structure kernels at C2
size (CSR build, degree bucketing, unique).  One JSON line per measurement -> gpurun_out/configs.jsonl.

Small configs are L2-resident and launch-latency-bound: the number of record is
microseconds per op; the roofline fraction is informational.  Timing: CUDA events on the launching
stream, 5 warm-ups, median of 20; small working sets stay in L2 on purpose (that IS their regime),
the C2-size structure kernels stream far more than the 126 MB L2.
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "dgl-1_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

from benchmarks import graphs  # noqa: E402
from dgl_b200 import ops  # noqa: E402

def _peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"])
    except Exception:  # noqa: BLE001
        return 6540.5


PEAK = _peak()
DEV = torch.device("cuda")


def timeit(fn, warm=5, iters=20):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


OUT = []


def report(config, op, ms, alg_bytes, **extra):
    rec = dict(config=config, op=op, us=round(ms * 1e3, 2), alg_bytes=int(alg_bytes),
               alg_GBps=round(alg_bytes / max(ms, 1e-9) / 1e6, 1),
               frac_of_hbm_peak=round(alg_bytes / max(ms, 1e-9) / 1e6 / PEAK, 4))
    rec.update(extra)
    OUT.append(rec)
    print(json.dumps(rec), flush=True)


def csr_of(src, dst, n):
    return ops.csr_build(dst, src, n, n), ops.csr_build(src, dst, n, n)


def c1():
    c = graphs.C1_CORA
    n = c["n"]
    src, dst = graphs.uniform_graph(n, c["e"], c["seed"], DEV, self_loops=True)
    cin, cout = csr_of(src, dst, n)
    e = cin.nnz
    for feat in (c["hidden"], c["classes"], c["feat_in"]):
        x = torch.randn(n, feat, device=DEV)
        ms = timeit(lambda: ops.spmm(cin, ops.MSG_COPY_SRC, ops.RED_SUM, x=x))
        report("C1", "spmm copy_src->sum F=%d" % feat, ms, graphs.spmm_alg_bytes(e, n, feat), n=n, e=e)
        ms = timeit(lambda: ops.spmm_backward_x(cout, x))
        report("C1", "spmm gradient F=%d" % feat, ms, graphs.spmm_alg_bytes(e, n, feat), n=n, e=e)
    ms = timeit(lambda: csr_of(src, dst, n))
    report("C1", "csr_build in+out", ms, 2 * e * 24, n=n, e=e)


def c3():
    c = graphs.C3_PUBMED
    n, heads, hid = c["n"], c["heads"], c["hidden"]
    src, dst = graphs.uniform_graph(n, c["e"], c["seed"], DEV, self_loops=True)
    cin, cout = csr_of(src, dst, n)
    e = cin.nnz
    ft = torch.randn(n, heads, hid, device=DEV)
    a1 = torch.randn(n, heads, device=DEV)
    a2 = torch.randn(n, heads, device=DEV)
    fused_bytes = e * (heads * hid * 4 + heads * 4 + 4) + n * heads * (hid + 2) * 4
    ms = timeit(lambda: ops.gat_aggregate(cin, a1, a2, ft, 0.2, flavour=ops.SOFTMAX_GAT_EXAMPLE, want_edge=False))
    report("C3", "fused GAT aggregation (example flavour)", ms, fused_bytes, n=n, e=e)
    ms = timeit(lambda: ops.gat_aggregate(cin, a1, a2, ft, 0.2, flavour=ops.SOFTMAX_STABLE, want_edge=False))
    report("C3", "fused GAT aggregation (stable softmax)", ms, fused_bytes, n=n, e=e)

    def unfused():
        s = ops.sddmm(cin, ops.SDDMM_ADD, a1, a2, heads, 1)
        alpha = ops.edge_softmax(cin, s)
        return ops.spmm(cin, ops.MSG_SRC_MUL_EDGE, ops.RED_SUM, x=ft.reshape(n, -1), w=alpha, w_width=heads)
    ms = timeit(unfused)
    report("C3", "sddmm(add) + edge_softmax + spmm(src_mul_edge) [3 kernels]", ms, fused_bytes + 4 * e * heads * 4, n=n, e=e)
    u = torch.randn(n, heads * hid, device=DEV)
    ms = timeit(lambda: ops.sddmm(cin, ops.SDDMM_DOT, u, u, heads, hid))
    report("C3", "sddmm u.v H=8 D=8", ms, e * (2 * heads * hid * 4 + 4 + heads * 4), n=n, e=e)


def c4():
    c = graphs.C4_AIFB
    n, hid = c["n"], c["hidden"]
    src, dst = graphs.uniform_graph(n, c["e"], c["seed"], DEV)
    cin, _ = csr_of(src, dst, n)
    e = cin.nnz
    x = torch.randn(n, hid, device=DEV)
    w = torch.rand(e, device=DEV)
    ms = timeit(lambda: ops.spmm(cin, ops.MSG_SRC_MUL_EDGE, ops.RED_SUM, x=x, w=w.reshape(-1, 1), w_width=1))
    report("C4", "spmm src_mul_edge(scalar norm)->sum F=16, all relations in one launch", ms,
           graphs.spmm_alg_bytes(e, n, hid, weights=True), n=n, e=e)
    m = torch.randn(e, hid, device=DEV)
    ms = timeit(lambda: ops.spmm(cin, ops.MSG_COPY_EDGE, ops.RED_SUM, w=m))
    report("C4", "spmm copy_edge->sum F=16 (typed messages)", ms, e * (hid * 4 + 4) + n * hid * 4, n=n, e=e)
    ms = timeit(lambda: ops.spmm(cin, ops.MSG_COPY_SRC, ops.RED_MAX, x=x, want_arg=True))
    report("C4", "spmm copy_src->max + argmax F=16", ms, graphs.spmm_alg_bytes(e, n, hid) + n * hid * 4, n=n, e=e)


def c4_per_relation():
    """C4 also as 90 per-relation src_mul_edge->sum launches (the reference's R-GCN loops
    over relations in Python: examples/pytorch/rgcn/layers.py) against the one batched launch."""
    c = graphs.C4_AIFB
    n, hid, rels = c["n"], c["hidden"], c["rels"]
    src, dst = graphs.uniform_graph(n, c["e"], c["seed"], DEV)
    e = src.shape[0]
    g = torch.Generator(device=DEV)
    g.manual_seed(c["seed"])
    rel = torch.randint(0, rels, (e,), device=DEV, generator=g)
    x = torch.randn(n, hid, device=DEV)
    w = torch.rand(e, 1, device=DEV)
    subs = []
    for r in range(rels):
        m = rel == r
        subs.append((ops.csr_build(dst[m], src[m], n, n), w[m].contiguous()))

    def per_relation():
        acc = torch.zeros(n, hid, device=DEV)
        for csr, wr in subs:
            acc += ops.spmm(csr, ops.MSG_SRC_MUL_EDGE, ops.RED_SUM, x=x, w=wr, w_width=1)[0]
        return acc
    ms = timeit(per_relation)
    report("C4", "90 per-relation src_mul_edge->sum launches + accumulate", ms,
           graphs.spmm_alg_bytes(e, n, hid, weights=True) + 2 * rels * n * hid * 4, n=n, e=e)
    cin = ops.csr_build(dst, src, n, n)
    ms = timeit(lambda: ops.spmm(cin, ops.MSG_SRC_MUL_EDGE, ops.RED_SUM, x=x, w=w, w_width=1))
    report("C4", "the same as ONE batched launch", ms, graphs.spmm_alg_bytes(e, n, hid, weights=True), n=n, e=e)
    got = ops.spmm(cin, ops.MSG_SRC_MUL_EDGE, ops.RED_SUM, x=x, w=w, w_width=1)[0]
    want = per_relation()
    assert torch.allclose(got, want, rtol=1e-4, atol=1e-4), "batched != per-relation"


def c2_torch_sparse():
    """'Reference on GPU' row (optional): the reference's own GPU path is th.sparse.mm on
    the uncoalesced COO it builds in edge-id order (python/dgl/backend/pytorch/tensor.py:45-51,146).
    Library call timed beside the engine on the same graph and features; forward and fwd+bwd."""
    c = graphs.C2_REDDIT
    n, e = c["n"], c["e"]
    src, dst = graphs.powerlaw_symmetric(n, e, c["seed"], DEV)
    idx = torch.stack([dst, src])
    del src, dst
    adj = torch.sparse_coo_tensor(idx, torch.ones(e, device=DEV), (n, n))
    x = torch.randn(n, 256, device=DEV, requires_grad=True)
    dy = torch.randn(n, 256, device=DEV)
    alg = graphs.spmm_alg_bytes(e, n, 256)
    try:
        ms = timeit(lambda: torch.sparse.mm(adj, x.detach()), warm=2, iters=5)
        report("C2", "REFERENCE ON GPU: th.sparse.mm(uncoalesced COO) forward F=256", ms, alg, n=n, e=e)

        def fb():
            y = torch.sparse.mm(adj, x)
            y.backward(dy)
            x.grad = None
        ms = timeit(fb, warm=2, iters=5)
        report("C2", "REFERENCE ON GPU: th.sparse.mm fwd + autograd bwd F=256", ms, 2 * alg, n=n, e=e)
        adj_c = adj.coalesce()
        ms = timeit(lambda: torch.sparse.mm(adj_c, x.detach()), warm=2, iters=5)
        report("C2", "th.sparse.mm(pre-coalesced COO) forward F=256", ms, alg, n=n, e=e)
    except RuntimeError as err:      # e.g. out of memory inside the library's coalesce
        report("C2", "REFERENCE ON GPU: th.sparse.mm failed: %s" % str(err)[:80], float("nan"), 0, n=n, e=e)


def c2_structure():
    c = graphs.C2_REDDIT
    n, e = c["n"], c["e"]
    src, dst = graphs.powerlaw_symmetric(n, e, c["seed"], DEV)
    ms = timeit(lambda: ops.csr_build(dst, src, n, n), warm=2, iters=5)
    report("C2", "csr_build (in-edge CSR from 114.6M-edge COO)", ms, e * (16 + 8) + 3 * 3 * e * 8, n=n, e=e)
    mids = torch.arange(e, device=DEV)
    recv = torch.arange(n, device=DEV)
    ms = timeit(lambda: ops.degree_bucketing(mids, dst, recv, n), warm=1, iters=3)
    report("C2", "degree_bucketing (114.6M messages, 233K nodes)", ms, e * 8 * 6, n=n, e=e)
    ms = timeit(lambda: ops.unique_sorted(dst, n), warm=1, iters=3)
    report("C2", "unique_sorted + inverse (114.6M ids)", ms, e * 8 * 5, n=n, e=e)
    cin = ops.csr_build(dst, src, n, n)
    del src, dst
    for dt, b in ((torch.float32, 4), (torch.bfloat16, 2)):
        x = torch.randn(n, 256, device=DEV).to(dt)
        ms = timeit(lambda: ops.spmm(cin, ops.MSG_COPY_SRC, ops.RED_SUM, x=x), warm=3, iters=10)
        report("C2", "spmm copy_src->sum F=256 %s" % str(dt).split(".")[-1], ms,
               graphs.spmm_alg_bytes(e, n, 256, elem=b), n=n, e=e)
    x = torch.randn(n, 256, device=DEV)
    w = torch.rand(e, 1, device=DEV)
    ms = timeit(lambda: ops.spmm(cin, ops.MSG_SRC_MUL_EDGE, ops.RED_SUM, x=x, w=w, w_width=1), warm=3, iters=10)
    report("C2", "spmm src_mul_edge(scalar)->sum F=256 f32", ms, graphs.spmm_alg_bytes(e, n, 256, weights=True), n=n, e=e)
    wh = torch.rand(e, 8, device=DEV)
    ms = timeit(lambda: ops.spmm(cin, ops.MSG_SRC_MUL_EDGE, ops.RED_SUM, x=x, w=wh, w_width=8), warm=3, iters=10)
    report("C2", "spmm src_mul_edge(8 heads x 32)->sum F=256 f32", ms,
           graphs.spmm_alg_bytes(e, n, 256) + e * (8 * 4 + 4), n=n, e=e)
    del wh
    ms = timeit(lambda: ops.edge_data_to_slots(cin, w), warm=3, iters=10)
    report("C2", "edge_data_to_slots (w[eid] -> slot order)", ms, e * 12, n=n, e=e)
    ms = timeit(lambda: ops.spmm(cin, ops.MSG_COPY_SRC, ops.RED_MAX, x=x, want_arg=True), warm=3, iters=10)
    report("C2", "spmm copy_src->max+argmax F=256 f32", ms, graphs.spmm_alg_bytes(e, n, 256) + n * 256 * 4, n=n, e=e)
    for feat in (602, 41):
        x = torch.randn(n, feat, device=DEV)
        ms = timeit(lambda: ops.spmm(cin, ops.MSG_COPY_SRC, ops.RED_SUM, x=x), warm=2, iters=5)
        report("C2", "spmm copy_src->sum F=%d f32" % feat, ms, graphs.spmm_alg_bytes(e, n, feat), n=n, e=e)


def c3_at_c2_size():
    """Kernel family (3) at scale (round-1 review: no measurement of any kind): the C2 graph (114.6M
    edges, hubs of degree 22K) with 8 heads x 32 = 256 features - SDDMM add / dot, edge softmax forward
    and backward, the fused GAT aggregation in both flavours."""
    c = graphs.C2_REDDIT
    n, e, heads, hid = c["n"], c["e"], 8, 32
    src, dst = graphs.powerlaw_symmetric(n, e, c["seed"], DEV)
    cin = ops.csr_build(dst, src, n, n)
    del src, dst
    a1 = torch.randn(n, heads, device=DEV)
    a2 = torch.randn(n, heads, device=DEV)
    ft = torch.randn(n, heads, hid, device=DEV)
    ms = timeit(lambda: ops.sddmm(cin, ops.SDDMM_ADD, a1, a2, heads, 1), warm=2, iters=5)
    report("C2x8h", "sddmm u+v H=8 D=1 (GAT scores)", ms, e * (2 * heads * 4 + 4 + heads * 4), n=n, e=e)
    s = ops.sddmm(cin, ops.SDDMM_ADD, a1, a2, heads, 1)
    ms = timeit(lambda: ops.edge_softmax(cin, s), warm=2, iters=5)
    report("C2x8h", "edge_softmax H=8", ms, e * heads * 4 * 2 + e * 4, n=n, e=e)
    alpha = ops.edge_softmax(cin, s)
    ms = timeit(lambda: ops.edge_softmax_backward(cin, alpha, s), warm=2, iters=5)
    report("C2x8h", "edge_softmax backward H=8", ms, e * heads * 4 * 3 + e * 4, n=n, e=e)
    del s, alpha
    # the same three kernels with edge data in CSR SLOT order (eids = NULL: what Adjacency.slot_view() hands them)
    cslot = ops.CSR(n, n, cin.indptr, cin.indices, None)
    ms = timeit(lambda: ops.sddmm(cslot, ops.SDDMM_ADD, a1, a2, heads, 1), warm=2, iters=5)
    report("C2x8h", "sddmm u+v H=8 D=1, slot order", ms, e * (2 * heads * 4 + 4 + heads * 4), n=n, e=e)
    s = ops.sddmm(cslot, ops.SDDMM_ADD, a1, a2, heads, 1)
    ms = timeit(lambda: ops.edge_softmax(cslot, s), warm=2, iters=5)
    report("C2x8h", "edge_softmax H=8, slot order", ms, e * heads * 4 * 2, n=n, e=e)
    alpha = ops.edge_softmax(cslot, s)
    ms = timeit(lambda: ops.edge_softmax_backward(cslot, alpha, s), warm=2, iters=5)
    report("C2x8h", "edge_softmax backward H=8, slot order", ms, e * heads * 4 * 3, n=n, e=e)
    del s, alpha
    u = ft.reshape(n, -1)
    ms = timeit(lambda: ops.sddmm(cin, ops.SDDMM_DOT, u, u, heads, hid), warm=2, iters=5)
    report("C2x8h", "sddmm u.v H=8 D=32", ms, e * (2 * heads * hid * 4 + 4 + heads * 4), n=n, e=e)
    fused_bytes = e * (heads * hid * 4 + heads * 4 + 4) + n * heads * (hid + 2) * 4
    for fl, name in ((ops.SOFTMAX_GAT_EXAMPLE, "example"), (ops.SOFTMAX_STABLE, "stable")):
        ms = timeit(lambda: ops.gat_aggregate(cin, a1, a2, ft, 0.2, flavour=fl, want_edge=False), warm=1, iters=3)
        report("C2x8h", "fused GAT aggregation (%s flavour) H=8 D=32" % name, ms, fused_bytes, n=n, e=e)


def gat_chain():
    """The differentiable GAT edge path (edge_ops.gat_attention: SDDMM -> leaky_relu -> edge softmax -> weighted SpMM,
    forward + backward) at C2 size with 8 heads x 16, through the edge-id indirection and over the slot view."""
    from dgl_b200 import autograd, edge_ops
    c = graphs.C2_REDDIT
    n, e, heads, hid = c["n"], c["e"], 8, 16
    src, dst = graphs.powerlaw_symmetric(n, e, c["seed"], DEV)
    adj = autograd.Adjacency(n, n, coo=(dst, src, None))
    _ = adj.csr_in, adj.csr_out
    adj.slot_view()
    adj.drop_coo()
    del src, dst
    torch.cuda.empty_cache()
    a1 = torch.randn(n, heads, device=DEV, requires_grad=True)
    a2 = torch.randn(n, heads, device=DEV, requires_grad=True)
    ft = torch.randn(n, heads, hid, device=DEV, requires_grad=True)
    gout = torch.randn(n, heads, hid, device=DEV)
    for slot in (False, True):
        def step():
            out, _ = edge_ops.gat_attention(adj, a1, a2, ft, 0.2, want_alpha=False, slot_order=slot)
            out.backward(gout)
            a1.grad = a2.grad = ft.grad = None
        ms = timeit(step, warm=2, iters=5)
        report("C2x8h", "gat_attention forward + backward (autograd chain), %s" % ("slot order" if slot else "edge-id order"),
               ms, 0, n=n, e=e, heads=heads, dim=hid)


def c5_kernel():
    """C5 on one GPU: the SpMM kernel alone (N = 50M, E = 150M, F = 128), fp32 and bf16."""
    c = graphs.C5_RMAT
    n, e, feat = c["n"], c["e"], c["feat"]
    src, dst = graphs.rmat(26, e, c["seed"], DEV, n=n)
    cin = ops.csr_build(dst, src, n, n)
    del src, dst
    for dt, b in ((torch.float32, 4), (torch.bfloat16, 2)):
        x = torch.empty(n, feat, device=DEV, dtype=dt)
        for s0 in range(0, n, 1 << 22):
            x[s0:s0 + (1 << 22)] = torch.randn(min(1 << 22, n - s0), feat, device=DEV).to(dt)
        y = torch.empty(n, feat, device=DEV, dtype=dt)
        ms = timeit(lambda: ops.spmm(cin, ops.MSG_COPY_SRC, ops.RED_SUM, x=x, out=y), warm=2, iters=6)
        report("C5", "spmm copy_src->sum F=128 %s" % str(dt).split(".")[-1], ms,
               graphs.spmm_alg_bytes(e, n, feat, elem=b), n=n, e=e,
               min_bytes=graphs.spmm_min_bytes(e, n, n, feat, elem=b))
        del x, y


def cpu_layers():
    """CPU baselines of the GAT and R-GCN example layers: the arithmetic the reference's
    examples run per layer (examples/pytorch/gat/train.py:61-97 - gathers, leaky_relu, exp().clamp,
    two incidence SpMMs, divide; examples/pytorch/rgcn/layers.py:86-100 - per-relation
    src_mul_edge -> sum = th.sparse.mm per relation), with torch on the host cores."""
    import time
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    c = graphs.C3_PUBMED
    n, heads, hid = c["n"], c["heads"], c["hidden"]
    src, dst = graphs.uniform_graph(n, c["e"], c["seed"], "cpu", self_loops=True)
    e = src.shape[0]
    ft = torch.randn(n, heads, hid)
    a1, a2 = torch.randn(n, heads, 1), torch.randn(n, heads, 1)
    inc = torch.sparse_coo_tensor(torch.stack([dst, torch.arange(e)]), torch.ones(e), (n, e))

    def gat_layer():
        a = torch.nn.functional.leaky_relu(a1[src] + a2[dst], 0.2).exp().clamp(-10, 10)      # (E, H, 1)
        m = ft[src] * a                                                                        # (E, H, D)
        num = torch.sparse.mm(inc, m.reshape(e, -1)).reshape(n, heads, hid)
        z = torch.sparse.mm(inc, a.reshape(e, -1)).reshape(n, heads, 1)
        return num / z
    gat_layer()
    ts = []
    for _ in range(5):
        t0 = time.perf_counter()
        gat_layer()
        ts.append(time.perf_counter() - t0)
    fused_bytes = e * (heads * hid * 4 + heads * 4 + 4) + n * heads * (hid + 2) * 4
    report("C3", "CPU BASELINE: GAT layer edge path (torch, %d threads)" % torch.get_num_threads(),
           sorted(ts)[2] * 1e3, fused_bytes, n=n, e=e)
    c = graphs.C4_AIFB
    n, hid, rels = c["n"], c["hidden"], c["rels"]
    src, dst = graphs.uniform_graph(n, c["e"], c["seed"], "cpu")
    e = src.shape[0]
    g = torch.Generator().manual_seed(c["seed"])
    rel = torch.randint(0, rels, (e,), generator=g)
    x = torch.randn(n, hid)
    w = torch.rand(e)
    adjs = []
    for r in range(rels):
        m = rel == r
        adjs.append(torch.sparse_coo_tensor(torch.stack([dst[m], src[m]]), w[m], (n, n)))

    def rgcn_layer():
        acc = torch.zeros(n, hid)
        for a in adjs:
            acc += torch.sparse.mm(a, x)
        return acc
    rgcn_layer()
    ts = []
    for _ in range(5):
        t0 = time.perf_counter()
        rgcn_layer()
        ts.append(time.perf_counter() - t0)
    report("C4", "CPU BASELINE: R-GCN layer, 90 per-relation th.sparse.mm (torch, %d threads)" % torch.get_num_threads(),
           sorted(ts)[2] * 1e3, graphs.spmm_alg_bytes(e, n, hid, weights=True), n=n, e=e)


def sampling():
    """Neighbour sampling on a C2-shaped graph (232,965 nodes, power-law degrees; 1/5 of
    Reddit's edges so that the reference's ImmutableGraph - two std::sorts of the edge list on one core - builds
    in a bounded time): batches of 1000 seeds, fanout 10, 2 hops, in-neighbours.
      * this engine: dgl_b200.sampling.sample_subgraph end to end (per hop: offsets + the sampling kernel that
        also discovers the next frontier; then the bitmap compaction and the placement of the picks; 3 syncs that
        return sizes), and the sampling kernel alone over every row;
      * CPU arm: the UNMODIFIED compiled reference (oracle/_ref/libdgl.so) through its C ABI - one batch per call
        (one core) and 16 batches per call (its OpenMP path, graph_apis.cc:451-454)."""
    import time
    import numpy as np
    from dgl_b200 import sampling as engine
    c = graphs.C2_REDDIT
    n, e = c["n"], c["e"] // 5
    src, dst = graphs.powerlaw_symmetric(n, e, c["seed"], DEV)
    e = int(src.shape[0])
    csr = ops.csr_build(dst, src, n, n, order=1)
    gen = torch.Generator(device="cpu").manual_seed(0)
    batches = [torch.randperm(n, generator=gen)[:1000].to(DEV) for _ in range(16)]
    fanout, hops = 10, 2
    state = {"i": 0, "nodes": 0, "edges": 0, "calls": 0}

    def one_batch():
        sub = engine.sample_subgraph(csr, batches[state["i"] % 16], fanout, hops, seed=state["i"])
        state["i"] += 1
        state["nodes"] += int(sub.induced_nodes.shape[0])
        state["edges"] += int(sub.induced_edges.shape[0])
        state["calls"] += 1
    ms = timeit(one_batch, warm=5, iters=40)
    nodes, edges = state["nodes"] / state["calls"], state["edges"] / state["calls"]
    report("C2/5", "neighbour sampling, 1000 seeds x fanout 10 x 2 hops: sample_subgraph end to end", ms,
           edges * (8 + 3 * 8), n=n, e=e, nodes_per_subgraph=round(nodes), edges_per_subgraph=round(edges),
           sampled_edges_per_s=round(edges / ms * 1e3))
    rows = torch.arange(n, dtype=torch.int64, device=DEV)
    for k in (10, 25):
        picks = int(torch.clamp(ops.degrees(csr), max=k).sum())
        ms_k = timeit(lambda: ops.neighbor_sample(csr, rows, k, seed=1), warm=3, iters=20)
        report("C2/5", "sampling kernel pair (offsets + fill) over all %d rows, fanout %d" % (n, k), ms_k,
               n * 32 + picks * (8 + 3 * 8), picks=picks, sampled_edges_per_s=round(picks / ms_k * 1e3))
    del csr
    # ---- CPU arm: the compiled reference
    from oracle import ref_ffi
    if not ref_ffi.available():
        report("C2/5", "CPU BASELINE: reference NeighborUniformSample - oracle/_ref/libdgl.so not present", 0.0, 0)
        return
    t0 = time.perf_counter()
    g = ref_ffi.RefGraph(src.cpu().numpy(), dst.cpu().numpy(), n, readonly=True, multigraph=True)
    build_s = time.perf_counter() - t0
    seeds_np = [b.cpu().numpy() for b in batches]
    for per_call in (1, 16):
        ts, got_edges = [], 0
        for rep in range(6):
            t0 = time.perf_counter()
            res = ref_ffi.graph_uniform_sampling(g, seeds_np[:per_call], "in", hops, fanout)
            ts.append(time.perf_counter() - t0)
            got_edges = sum(len(r[1]) for r in res) / per_call
        t = sorted(ts[1:])[len(ts[1:]) // 2] / per_call
        report("C2/5", "CPU BASELINE: reference NeighborUniformSample (libdgl.so, %d batch(es) per call, %d cores)"
               % (per_call, min(per_call, os.cpu_count() or 1)), t * 1e3, got_edges * (8 + 3 * 8),
               edges_per_subgraph=round(got_edges), sampled_edges_per_s=round(got_edges / t),
               immutable_graph_build_s=round(build_s, 1), speedup_of_engine=round(t * 1e3 / ms, 2))


def _capture(step, warm=3):
    """torch.cuda.CUDAGraph of `step` (which must use static input tensors): warm up on a side stream, capture,
    return the graph.  Any cudaStreamSynchronize / size-returning call inside `step` makes the capture fail -
    which is the point of the exercise (round-1 review: "impossible as written")."""
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(warm):
            step()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    gph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gph):
        out = step()
    return gph, out


def graph_capture():
    """The launch-latency-bound configs as ONE CUDA graph per training step: the structure (CSR
    pair, long-row plan) is built once per graph, the steady-state step of the builtin path has no
    size-returning call left in it, so a whole forward + backward + SGD update is captured and replayed.
      C1: 2-layer GCN (gcn_spmv.py) on the Cora-shaped graph through DGLGraph.update_all;
      C3: one GAT layer (8 heads x 8) on the PubMed-shaped graph through gat_attention (SDDMM -> edge softmax ->
          weighted SpMM and their gradient kernels);
      C4: one R-GCN layer on the AIFB-shaped graph, src_mul_edge -> sum with a per-edge norm, all relations in one
          launch."""
    import dgl_b200
    import dgl_b200.function as fn
    from dgl_b200 import _lib, autograd, edge_ops
    lr = 1e-2

    def sgd(params):
        with torch.no_grad():
            for p in params:
                p.add_(p.grad, alpha=-lr)
                p.grad = None

    only = os.environ.get("CAPTURE_ONLY")          # a failed capture poisons the process: one config per process

    def measure(tag, step, params, n, e):
        # capture FIRST: the parameters' AccumulateGrad nodes must not have been created on the default stream
        try:
            gph, out = _capture(step)
        except Exception as ex:  # noqa: BLE001
            import traceback
            traceback.print_exc()
            report(tag, "training step: CUDA-graph capture FAILED: %s | %s" % (str(ex).splitlines()[0][:200],
                                                                              _lib.last_error()[:200]), 0.0, 0, n=n, e=e)
            return
        l0 = _lib.launch_count()
        step()
        torch.cuda.synchronize()
        launches = _lib.launch_count() - l0
        eager_ms = timeit(step, warm=5, iters=30)
        # replay equals eager: same weights in, same loss / weights out
        saved = [p.detach().clone() for p in params]
        gph.replay()
        torch.cuda.synchronize()
        after_graph = [p.detach().clone() for p in params]
        with torch.no_grad():
            for p, q in zip(params, saved):
                p.copy_(q)
        torch.manual_seed(0)
        step()
        torch.cuda.synchronize()
        same = all(torch.allclose(a, p.detach(), rtol=1e-5, atol=1e-6) for a, p in zip(after_graph, params))
        graph_ms = timeit(gph.replay, warm=5, iters=30)
        report(tag, "training step (fwd + bwd + SGD): eager", eager_ms, 0, n=n, e=e, engine_launches_per_step=launches)
        report(tag, "training step (fwd + bwd + SGD): ONE CUDA graph replay", graph_ms, 0, n=n, e=e,
               engine_launches_per_step=launches, speedup_vs_eager=round(eager_ms / graph_ms, 2),
               replay_equals_eager=bool(same))

    # ---- C1
    c = graphs.C1_CORA
    n = c["n"]
    src, dst = graphs.uniform_graph(n, c["e"], c["seed"], DEV, self_loops=True)
    g = dgl_b200.DGLGraph()
    g.add_nodes(n)
    g.add_edges(src, dst)
    g.to(DEV)
    norm = g.in_degrees().float().clamp(min=1).pow(-0.5).unsqueeze(1)
    x = torch.randn(n, c["feat_in"], device=DEV)
    labels = torch.randint(0, c["classes"], (n,), device=DEV)
    w1 = (torch.randn(c["feat_in"], c["hidden"], device=DEV) * 0.05).requires_grad_(True)
    w2 = (torch.randn(c["hidden"], c["classes"], device=DEV) * 0.05).requires_grad_(True)
    b1 = torch.zeros(c["hidden"], device=DEV, requires_grad=True)
    b2 = torch.zeros(c["classes"], device=DEV, requires_grad=True)

    def agg(h):
        g.ndata["h"] = h
        g.update_all(fn.copy_src("h", "m"), fn.sum("m", "h"))
        return g.ndata.pop("h")

    def gcn_step():
        h = torch.relu(agg(torch.mm(x, w1) * norm) * norm + b1)
        logits = agg(torch.mm(h, w2) * norm) * norm + b2
        loss = torch.nn.functional.cross_entropy(logits, labels)
        loss.backward()
        sgd([w1, w2, b1, b2])
        return loss
    if only in (None, "C1"):
        g.ndata["h"] = torch.zeros(n, 1, device=DEV)
        g.update_all(fn.copy_src("h", "m"), fn.sum("m", "h"))     # structure + plan built before the capture
        measure("C1", gcn_step, [w1, w2, b1, b2], n, int(src.shape[0]))

    # ---- C3
    c = graphs.C3_PUBMED
    n, heads, hid = c["n"], c["heads"], c["hidden"]
    src, dst = graphs.uniform_graph(n, c["e"], c["seed"], DEV, self_loops=True)
    adj = autograd.Adjacency(n, n, coo=(dst, src, None))
    _ = adj.csr_in, adj.csr_out
    x3 = torch.randn(n, c["feat_in"], device=DEV)
    wf = (torch.randn(c["feat_in"], heads * hid, device=DEV) * 0.05).requires_grad_(True)
    al = (torch.randn(heads, hid, device=DEV) * 0.1).requires_grad_(True)
    ar = (torch.randn(heads, hid, device=DEV) * 0.1).requires_grad_(True)

    def gat_step():
        ft = torch.mm(x3, wf).reshape(n, heads, hid)
        a1 = (ft * al).sum(-1)
        a2 = (ft * ar).sum(-1)
        out, _ = edge_ops.gat_attention(adj, a1, a2, ft, 0.2)
        loss = out.square().mean()
        loss.backward()
        sgd([wf, al, ar])
        return loss
    if only in (None, "C3"):
        measure("C3", gat_step, [wf, al, ar], n, int(src.shape[0]))

    # ---- C4
    c = graphs.C4_AIFB
    n, hid = c["n"], c["hidden"]
    src, dst = graphs.uniform_graph(n, c["e"], c["seed"], DEV)
    adj4 = autograd.Adjacency(n, n, coo=(dst, src, None))
    _ = adj4.csr_in, adj4.csr_out
    e4 = int(src.shape[0])
    enorm = torch.rand(e4, 1, device=DEV)
    emb = (torch.randn(n, hid, device=DEV) * 0.1).requires_grad_(True)
    w4 = (torch.randn(hid, hid, device=DEV) * 0.1).requires_grad_(True)

    def rgcn_step():
        h = autograd.aggregate(adj4, ops.MSG_SRC_MUL_EDGE, ops.RED_SUM, x=torch.mm(emb, w4), w=enorm, w_width=1)
        loss = torch.relu(h).square().mean()
        loss.backward()
        sgd([emb, w4])
        return loss
    if only in (None, "C4"):
        measure("C4", rgcn_step, [emb, w4], n, e4)


if __name__ == "__main__":
    sections = dict(c1=c1, c3=c3, c4=c4, c4rel=c4_per_relation, c2=c2_structure, c2ref=c2_torch_sparse,
                    c3big=c3_at_c2_size, c5=c5_kernel, cpu=cpu_layers, sampling=sampling, capture=graph_capture, gatchain=gat_chain)
    for name in (sys.argv[1:] or list(sections)):
        sections[name]()
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", os.environ.get("CONFIGS_OUT", "configs.jsonl")), "w") as f:
        for r in OUT:
            f.write(json.dumps(r) + "\n")
