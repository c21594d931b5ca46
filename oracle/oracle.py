"""CPU restatement (numpy) of the reference's builtin message-passing path.

TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is imported by the product package
(dgl-1_b200/); only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs use it, and only as the checker / the timed CPU arm.

Parity status: PINNED.  tests/test_oracle_golden.py checks every function here against
  (a) the literal known-answer vectors of the reference's tests/compute/test_function.py:24-49,
  (b) golden vectors produced by the UNMODIFIED reference (C++ built by oracle/Makefile into
      oracle/_ref/libdgl.so + /root/reference/python imported in the build container) by
      oracle/gen_golden.py and committed under tests/golden/,
  (c) on machines where oracle/_ref/libdgl.so exists, the reference's own C++ GetAdj /
      DegreeBucketing called live through oracle/ref_ffi.py.

Every function cites the reference code it restates (paths relative to /root/reference).
All ids are int64, features float32, like the reference.
"""
import numpy as np


# ---------------------------------------------------------------------------------------
# graph structure
# ---------------------------------------------------------------------------------------

def get_adj_coo(src, dst, transpose=False):
    """Graph::GetAdj(transpose, "coo"), src/graph/graph.cc:509-524:
    idx = [dst | src] in edge-id order (or [src | dst] when transposed) + identity eid."""
    src = np.asarray(src, dtype=np.int64)
    dst = np.asarray(dst, dtype=np.int64)
    idx = np.concatenate([src, dst]) if transpose else np.concatenate([dst, src])
    return idx, np.arange(len(src), dtype=np.int64)


def get_adj_csr(src, dst, n, transpose=False):
    """Graph::GetAdj(transpose, "csr"), src/graph/graph.cc:525-549: concatenation of the
    per-node adjacency lists (in-edges unless transposed); a list holds its edges in insertion
    order, i.e. ascending edge id (Graph::AddEdge, graph.cc:48-63).  Returns int64
    (indptr[n+1], indices[E], eid[E])."""
    src = np.asarray(src, dtype=np.int64)
    dst = np.asarray(dst, dtype=np.int64)
    key, other = (src, dst) if transpose else (dst, src)
    order = np.argsort(key, kind="stable")
    indptr = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(np.bincount(key, minlength=n), out=indptr[1:])
    return indptr, other[order].astype(np.int64), order.astype(np.int64)


def immutable_csr(src, dst, n, sort_on, eid=None):
    """ImmutableGraph::CSR::FromEdges(edges, sort_on, n), src/graph/immutable_graph.cc:206-237:
    sort by (end_points[sort_on], end_points[other]); indices = other endpoint, edge_ids
    carried.  std::sort leaves the order of exact duplicate (src, dst) pairs unspecified; this
    restatement (like the CUDA build) breaks those ties by ascending input position.
    sort_on: 0 = by src (out-CSR), 1 = by dst (in-CSR)."""
    src = np.asarray(src, dtype=np.int64)
    dst = np.asarray(dst, dtype=np.int64)
    ends = (src, dst)
    key, other = ends[sort_on], ends[1 - sort_on]
    order = np.lexsort((np.arange(len(key)), other, key))
    indptr = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(np.bincount(key, minlength=n), out=indptr[1:])
    eid = np.arange(len(key), dtype=np.int64) if eid is None else np.asarray(eid, dtype=np.int64)
    return indptr, other[order].astype(np.int64), eid[order]


def in_degrees(dst, n):
    """Graph::InDegrees, src/graph/graph.cc:412-424."""
    return np.bincount(np.asarray(dst, dtype=np.int64), minlength=n).astype(np.int64)


def unique_sorted(ids):
    """recv_nodes = F.sort_1d(F.unique(v)) + utils.build_relabel_map
    (python/dgl/runtime/scheduler.py:136,341; python/dgl/utils.py:321-361): sorted unique ids
    and, per input element, its position in that list."""
    ids = np.asarray(ids, dtype=np.int64)
    uniq, inv = np.unique(ids, return_inverse=True)
    return uniq.astype(np.int64), inv.astype(np.int64)


# ---------------------------------------------------------------------------------------
# messages and reductions (semantic spec = the builtins' __call__)
# ---------------------------------------------------------------------------------------

def _flat(x, n):
    x = np.asarray(x)
    return x.reshape(n, -1)


def message(msg_op, src, x=None, w=None):
    """The UDF-equivalent __call__ of the builtins (python/dgl/function/message.py:75-91
    src_mul_edge with rank alignment by appending 1s, :128-135 copy_src, :173-180 copy_edge),
    evaluated for every edge.  Returns (E, ...) float32 messages."""
    src = np.asarray(src, dtype=np.int64)
    if msg_op == "copy_src":
        return np.asarray(x)[src]
    if msg_op == "copy_edge":
        return np.asarray(w)
    if msg_op == "src_mul_edge":
        s = np.asarray(x)[src]
        e = np.asarray(w)
        rank = max(s.ndim, e.ndim)
        s = s.reshape(s.shape + (1,) * (rank - s.ndim))
        e = e.reshape(e.shape + (1,) * (rank - e.ndim))
        return (s * e).astype(np.float32)
    raise ValueError(msg_op)


def reduce_sum(dst, msgs, n_out):
    """SPMVExecutor.run -> th.sparse.mm on the uncoalesced COO built in edge-id order
    (python/dgl/runtime/ir/executor.py:452-473; python/dgl/backend/pytorch/tensor.py:45-51,
    143-146): out[dst_e] += msg_e in float32, edges visited in ascending edge id
    (np.add.at is unbuffered and visits indices in order).  Rows with no edge are 0."""
    dst = np.asarray(dst, dtype=np.int64)
    msgs = np.asarray(msgs, dtype=np.float32)
    out = np.zeros((n_out,) + msgs.shape[1:], dtype=np.float32)
    np.add.at(out, dst, msgs)
    return out


def reduce_max(dst, msgs, n_out, eid=None):
    """fn.max through degree bucketing: F.max(mailbox, 1) per bucket
    (python/dgl/function/reducer.py:45-46; python/dgl/backend/pytorch/tensor.py:95-97), rows
    without messages filled by the zero initializer (python/dgl/runtime/degree_bucketing.py:72-78,
    python/dgl/init.py:47-61).  Also returns the argmax edge id (first maximum in ascending
    edge order; -1 for empty rows) which the reference does not expose - parity for `arg` is
    pinned by this restatement only."""
    dst = np.asarray(dst, dtype=np.int64)
    msgs = np.asarray(msgs, dtype=np.float32)
    E = len(dst)
    eid = np.arange(E, dtype=np.int64) if eid is None else np.asarray(eid, dtype=np.int64)
    flat = msgs.reshape(E, -1)
    out = np.full((n_out, flat.shape[1]), -np.inf, dtype=np.float32)
    arg = np.full((n_out, flat.shape[1]), -1, dtype=np.int64)
    for e in range(E):  # small cases only
        v = dst[e]
        # NaN propagates like torch.max over the mailbox: the first NaN message wins and stays
        better = (arg[v] < 0) | (flat[e] > out[v]) | (np.isnan(flat[e]) & ~np.isnan(out[v]))
        out[v] = np.where(better, flat[e], out[v])
        arg[v] = np.where(better, eid[e], arg[v])
    out[arg < 0] = 0.0
    return out.reshape((n_out,) + msgs.shape[1:]), arg.reshape((n_out,) + msgs.shape[1:])


def update_all(src, dst, n, msg_op, reduce_op, x=None, w=None):
    """DGLGraph.update_all(builtin, builtin) (python/dgl/graph.py:2398-2441) for one pair."""
    m = message(msg_op, src, x, w)
    if reduce_op == "sum":
        return reduce_sum(dst, m, n)
    if reduce_op == "max":
        return reduce_max(dst, m, n)[0]
    raise ValueError(reduce_op)


def spmm_grad(src, dst, n, dy, w=None):
    """Backward of th.sparse.mm wrt the dense operand (autograd of the call at
    python/dgl/backend/pytorch/tensor.py:146): dx = A^T dy, i.e. dx[src_e] += w_e * dy[dst_e]."""
    src = np.asarray(src, dtype=np.int64)
    g = np.asarray(dy, dtype=np.float32)[np.asarray(dst, dtype=np.int64)]
    if w is not None:
        wv = np.asarray(w, dtype=np.float32)
        rank = max(g.ndim, wv.ndim)
        g = g * wv.reshape(wv.shape + (1,) * (rank - wv.ndim))
    out = np.zeros((n,) + g.shape[1:], dtype=np.float32)
    np.add.at(out, src, g.astype(np.float32))
    return out


# ---------------------------------------------------------------------------------------
# GAT edge path
# ---------------------------------------------------------------------------------------

def sddmm(src, dst, u, v, op):
    """Per-edge u-op-v.  'add' is edge_attention's edges.src['a1'] + edges.dst['a2']
    (examples/pytorch/gat/train.py:90-92); 'dot' over the last dim is d(loss)/d(w) of
    SPMV_WITH_DATA."""
    a = np.asarray(u, dtype=np.float32)[np.asarray(src, dtype=np.int64)]
    b = np.asarray(v, dtype=np.float32)[np.asarray(dst, dtype=np.int64)]
    if op == "add":
        return a + b
    if op == "mul":
        return a * b
    if op == "dot":
        return (a * b).sum(-1)
    raise ValueError(op)


def leaky_relu(x, slope):
    return np.where(x >= 0, x, x * np.float32(slope)).astype(np.float32)


def gat_example(src, dst, n, a1, a2, ft, slope):
    """The reference GAT layer's edge path (examples/pytorch/gat/train.py:61-97), dropout off:
    a = clamp(exp(leaky_relu(a1[src] + a2[dst])), -10, 10); ft' = sum a * ft[src]; z = sum a;
    out = ft' / z.  a1, a2: (N, H, 1); ft: (N, H, D).  Returns (out, a, z)."""
    src = np.asarray(src, dtype=np.int64)
    dst = np.asarray(dst, dtype=np.int64)
    s = leaky_relu(np.asarray(a1, np.float32)[src] + np.asarray(a2, np.float32)[dst], slope)
    a = np.clip(np.exp(s), -10, 10).astype(np.float32)
    num = reduce_sum(dst, np.asarray(ft, np.float32)[src] * a, n)
    z = reduce_sum(dst, a, n)
    with np.errstate(divide="ignore", invalid="ignore"):
        out = num / z
    return out, a, z


def edge_softmax(dst, score, n):
    """Numerically stable per-destination softmax of edge scores (E, H) (not in the reference;
    the north-star's fused edge softmax).  alpha_e = exp(s_e - max_v) / sum_v."""
    dst = np.asarray(dst, dtype=np.int64)
    s = np.asarray(score, dtype=np.float32)
    mx = np.full((n,) + s.shape[1:], -np.inf, dtype=np.float32)
    np.maximum.at(mx, dst, s)
    ex = np.exp(s - mx[dst]).astype(np.float32)
    den = np.zeros((n,) + s.shape[1:], dtype=np.float32)
    np.add.at(den, dst, ex)
    return (ex / den[dst]).astype(np.float32)


# ---------------------------------------------------------------------------------------
# degree bucketing
# ---------------------------------------------------------------------------------------

def degree_bucketing(msg_ids, vids, recv_ids):
    """sched::DegreeBucketing (src/scheduler/scheduler.cc:13-93) in CANONICAL order.

    The reference iterates std::unordered_map / unordered_set, so bucket order and node order
    inside a bucket are implementation-defined; what is defined is: which nodes share a bucket,
    the per-node message order (= input order, scheduler.cc:22-25), the trailing zero-degree
    bucket made of recv_ids that got no message (:33-39, :79-85), and the section sizes.
    Canonical form: non-zero buckets by ascending degree, nodes ascending inside a bucket,
    zero-degree bucket last with ascending ids.
    Returns int64 arrays (degs, nids, nid_section, mids, mid_section)."""
    msg_ids = np.asarray(msg_ids, dtype=np.int64)
    vids = np.asarray(vids, dtype=np.int64)
    recv_ids = np.asarray(recv_ids, dtype=np.int64)
    in_edges = {}
    for m, v in zip(msg_ids.tolist(), vids.tolist()):
        in_edges.setdefault(v, []).append(m)
    bkt = {}
    for v, ms in in_edges.items():
        bkt.setdefault(len(ms), []).append(v)
    zero = sorted(set(r for r in recv_ids.tolist() if r not in in_edges))
    degs, nids, nsec, mids, msec = [], [], [], [], []
    for deg in sorted(bkt):
        nodes = sorted(bkt[deg])
        degs.append(deg)
        nsec.append(len(nodes))
        msec.append(deg * len(nodes))
        for v in nodes:
            nids.append(v)
            mids.extend(in_edges[v])
    if zero:
        degs.append(0)
        nsec.append(len(zero))
        nids.extend(zero)
    i64 = lambda a: np.asarray(a, dtype=np.int64)
    return i64(degs), i64(nids), i64(nsec), i64(mids), i64(msec)


def canonicalize_bucketing(degs, nids, nid_section, mids, mid_section):
    """Bring an implementation-ordered DegreeBucketing result (e.g. the reference's own output)
    into the canonical form of `degree_bucketing` so the two can be compared bit for bit."""
    degs = np.asarray(degs).tolist()
    nids = np.asarray(nids).tolist()
    nsec = np.asarray(nid_section).tolist()
    mids = np.asarray(mids).tolist()
    buckets = []
    npos = mpos = 0
    for b, deg in enumerate(degs):
        cnt = nsec[b]
        nodes = nids[npos:npos + cnt]
        npos += cnt
        per_node = []
        for _ in range(cnt):
            per_node.append(mids[mpos:mpos + deg])
            mpos += deg
        buckets.append((deg, nodes, per_node))
    nz = sorted([b for b in buckets if b[0] > 0], key=lambda b: b[0])
    z = [b for b in buckets if b[0] == 0]
    o_degs, o_nids, o_nsec, o_mids, o_msec = [], [], [], [], []
    for deg, nodes, per_node in nz:
        order = sorted(range(len(nodes)), key=lambda i: nodes[i])
        o_degs.append(deg)
        o_nsec.append(len(nodes))
        o_msec.append(deg * len(nodes))
        for i in order:
            o_nids.append(nodes[i])
            o_mids.extend(per_node[i])
    for deg, nodes, _ in z:
        o_degs.append(0)
        o_nsec.append(len(nodes))
        o_nids.extend(sorted(nodes))
    i64 = lambda a: np.asarray(a, dtype=np.int64)
    return i64(o_degs), i64(o_nids), i64(o_nsec), i64(o_mids), i64(o_msec)


# ---------------------------------------------------------------------------------------
# neighbour sampling
# ---------------------------------------------------------------------------------------

def neighbor_sample_subgraph(indptr, indices, eids, seeds, fanout, num_hops, rng=None):
    """ImmutableGraph::NeighborUniformSample = SampleSubgraph + CompactSubgraph
    (src/graph/immutable_graph.cc:772-1008) over one CSR of the immutable graph (in-CSR for
    neighbor_type "in", else the out-CSR, :780): a BFS queue over (vertex, layer); every queued vertex
    below the last layer contributes min(degree, fanout) slots of its row - the whole row
    (GetUniformSample :710-716) or `fanout` distinct slots in ascending slot order (:717-741; drawn with
    `rng`, a numpy RandomState - the reference uses rand_r seeded by time(), so only calls that never
    meet degree > fanout are reproducible).  Returns int64 arrays
      induced_nodes (ascending), layer_ids, induced_edges (parent edge ids, subgraph-edge order),
      sub_indptr, sub_indices (compacted ids) - the CSR of the subgraph whose edge ids are 0..E-1."""
    indptr = np.asarray(indptr, dtype=np.int64)
    indices = np.asarray(indices, dtype=np.int64)
    eids = np.asarray(eids, dtype=np.int64)
    layer = {}
    queue = []
    for s in np.asarray(seeds, dtype=np.int64).tolist():         # :787-793
        if s not in layer:
            layer[s] = 0
            queue.append(s)
    picked = {}                                                  # vertex -> (neighbours, edge ids)
    q = 0
    while q < len(queue):                                        # :806-861
        v = queue[q]
        q += 1
        if layer[v] >= num_hops:
            continue
        lo, hi = int(indptr[v]), int(indptr[v + 1])
        deg = hi - lo
        if deg <= fanout:
            slots = np.arange(lo, hi)
        else:
            if rng is None:
                raise ValueError("degree %d > fanout %d needs an rng" % (deg, fanout))
            slots = lo + np.sort(rng.choice(deg, int(fanout), replace=False))
        picked[v] = (indices[slots], eids[slots])
        for u in indices[slots].tolist():                        # :852-860
            if u not in layer:
                layer[u] = layer[v] + 1
                queue.append(u)
    nodes = np.array(sorted(layer), dtype=np.int64)              # :875-895
    layers = np.array([layer[v] for v in nodes.tolist()], dtype=np.int64)
    sub_indptr = np.zeros(len(nodes) + 1, dtype=np.int64)
    cols, edges = [], []
    for i, v in enumerate(nodes.tolist()):                       # :919-946
        if v in picked:
            cols.append(picked[v][0])
            edges.append(picked[v][1])
            sub_indptr[i + 1] = sub_indptr[i] + len(picked[v][0])
        else:
            sub_indptr[i + 1] = sub_indptr[i]
    cols = np.concatenate(cols) if cols else np.zeros(0, dtype=np.int64)
    edges = np.concatenate(edges) if edges else np.zeros(0, dtype=np.int64)
    sub_indices = np.searchsorted(nodes, cols).astype(np.int64)  # CompactSubgraph :963-983
    return nodes, layers, edges.astype(np.int64), sub_indptr, sub_indices


def check_neighbor_subgraph(indptr, indices, eids, seeds, fanout, num_hops,
                            induced_nodes, layer_ids, induced_edges, sub_rows, sub_cols):
    """Everything SampleSubgraph guarantees about its output whatever the random draws were, asserted.
    (indptr, indices, eids): the parent CSR that was sampled; sub_rows / sub_cols: per subgraph edge (in
    subgraph-edge-id order) the compacted id of the row vertex (the one whose neighbour list was sampled)
    and of the sampled neighbour.  Pinned on real outputs of the reference (tests/golden/sampling_*.npz).
      1. induced_nodes strictly ascending; layer 0 is exactly the set of seeds;
      2. subgraph edges are grouped by row vertex in ascending order (the sub-CSR, :919-946);
      3. edge i is a slot of its row vertex in the parent CSR: column = the neighbour, edge id =
         induced_edges[i]; within a row the slots are distinct and ascending (:720-741);
      4. a vertex below the last layer has min(degree, fanout) edges - its whole row when
         degree <= fanout (:710-716); a vertex of the last layer has none (:814-815);
      5. a non-seed vertex's layer is 1 + the smallest layer among the rows that picked it (BFS, :852-860)."""
    indptr = np.asarray(indptr, dtype=np.int64)
    indices = np.asarray(indices, dtype=np.int64)
    eids = np.asarray(eids, dtype=np.int64)
    nodes = np.asarray(induced_nodes, dtype=np.int64)
    layers = np.asarray(layer_ids, dtype=np.int64)
    edges = np.asarray(induced_edges, dtype=np.int64)
    rows = np.asarray(sub_rows, dtype=np.int64)
    cols = np.asarray(sub_cols, dtype=np.int64)
    assert len(nodes) == len(layers) and len(edges) == len(rows) == len(cols)
    assert np.all(np.diff(nodes) > 0), "induced_nodes not strictly ascending"
    seeds = np.unique(np.asarray(seeds, dtype=np.int64))
    assert np.array_equal(nodes[layers == 0], seeds), "layer 0 != the seeds"
    assert layers.min(initial=0) >= 0 and layers.max(initial=0) <= num_hops
    assert np.all(np.diff(rows) >= 0), "subgraph edges not grouped by ascending row vertex"
    slot_of_eid = np.full(int(eids.max(initial=-1)) + 1, -1, dtype=np.int64)
    slot_of_eid[eids] = np.arange(len(eids))
    slots = slot_of_eid[edges]
    assert np.all(slots >= 0)
    pr = nodes[rows]
    assert np.all((slots >= indptr[pr]) & (slots < indptr[pr + 1])), "edge is not a slot of its row"
    assert np.array_equal(indices[slots], nodes[cols]), "column does not match the parent edge"
    same = rows[1:] == rows[:-1]
    assert np.all(slots[1:][same] > slots[:-1][same]), "slots of a row not distinct ascending"
    cnt = np.bincount(rows, minlength=len(nodes))
    deg = indptr[nodes + 1] - indptr[nodes]
    want = np.where(layers < num_hops, np.minimum(deg, fanout), 0)
    assert np.array_equal(cnt, want), "edges per vertex != min(degree, fanout)"
    best = np.full(len(nodes), num_hops + 1, dtype=np.int64)
    np.minimum.at(best, cols, layers[rows] + 1)
    non_seed = layers > 0
    assert np.array_equal(best[non_seed], layers[non_seed]), "layer ids are not BFS levels"
    assert np.all(best[~non_seed] >= 1)
    return True
