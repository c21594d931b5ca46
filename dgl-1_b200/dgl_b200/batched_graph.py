"""Batched graphs and per-graph readout (python/dgl/batched_graph.py:18-331, 362-815).

`batch()` is a disjoint union: node and edge ids of graph i are shifted by the nodes / edges of the
graphs before it (gi.disjoint_union, src/graph/graph_op.cc), features are concatenated.  Message
passing on the batch is the ordinary path (one SpMM over the union).

Readout (`sum_nodes`, `mean_nodes`, `max_nodes`, and the `_edges` variants) is a segment reduction
over CONTIGUOUS id ranges, which is kernel (1) on a membership CSR: row g = graph g, its slots =
the ids of the graph's nodes (indptr = prefix sum of batch_num_nodes, indices = 0..N-1).  The
reference runs `unsorted_1d_segment_sum` for sum / mean and a PYTHON LOOP over the graphs for max
(batched_graph.py:751-762); here each is one launch, differentiable through the same autograd
Function as update_all (the transposed membership CSR has one slot per node).
"""
import torch

from . import ops
from .autograd import Adjacency, aggregate
from .base import ALL, DGLError, is_all
from .graph import DGLGraph

__all__ = ["BatchedDGLGraph", "batch", "unbatch", "sum_nodes", "sum_edges", "mean_nodes",
           "mean_edges", "max_nodes", "max_edges"]


def _init_attrs(graph_list, attrs, mode):
    """batched_graph.py:150-181: which fields are batched."""
    def items_and_fields(g):
        if mode == "node":
            return g.number_of_nodes(), set(g.ndata.keys())
        return g.number_of_edges(), set(g.edata.keys())
    if attrs is None:
        return []
    if is_all(attrs):
        attrs, ref_i = set(), None
        for i, g in enumerate(graph_list):
            num, fields = items_and_fields(g)
            if num > 0 and len(fields) > 0:
                attrs, ref_i = fields, i
                break
        if attrs:
            for i, g in enumerate(graph_list):
                num, fields = items_and_fields(g)
                if fields != attrs and num > 0:
                    raise ValueError("Expect graph {0} and {1} to have the same {2} attributes when "
                                     "{2}_attrs=ALL, got {3} and {4}.".format(ref_i, i, mode, attrs, fields))
        return sorted(attrs)
    if isinstance(attrs, str):
        return [attrs]
    return list(attrs)


class BatchedDGLGraph(DGLGraph):
    """Disjoint union of a list of graphs; read-only (batched_graph.py:18-274)."""

    def __init__(self, graph_list, node_attrs=ALL, edge_attrs=ALL):
        super(BatchedDGLGraph, self).__init__()
        node_attrs = _init_attrs(graph_list, node_attrs, "node")
        edge_attrs = _init_attrs(graph_list, edge_attrs, "edge")
        self._batch_num_nodes, self._batch_num_edges = [], []
        srcs, dsts, off = [], [], 0
        dev = None
        for g in graph_list:
            if isinstance(g, BatchedDGLGraph):       # a batch of batches flattens (:209-214)
                self._batch_num_nodes += g._batch_num_nodes
                self._batch_num_edges += g._batch_num_edges
            else:
                self._batch_num_nodes.append(g.number_of_nodes())
                self._batch_num_edges.append(g.number_of_edges())
            if g.number_of_edges() > 0:
                s, d = g.edges()
                srcs.append(s + off)
                dsts.append(d + off)
            if g._device is not None:
                dev = g._device
            off += g.number_of_nodes()
        DGLGraph.add_nodes(self, off)
        if srcs:
            tgt = dev if dev is not None else srcs[0].device
            DGLGraph.add_edges(self, torch.cat([s.to(tgt) for s in srcs]), torch.cat([d.to(tgt) for d in dsts]))
        if dev is not None:
            self.to(dev)
        for key in node_attrs:
            self.ndata[key] = torch.cat([g.ndata[key] for g in graph_list if g.number_of_nodes() > 0], 0)
        for key in edge_attrs:
            self.edata[key] = torch.cat([g.edata[key] for g in graph_list if g.number_of_edges() > 0], 0)
        self._segments = {}

    @property
    def batch_size(self):
        return len(self._batch_num_nodes)

    @property
    def batch_num_nodes(self):
        return self._batch_num_nodes

    @property
    def batch_num_edges(self):
        return self._batch_num_edges

    def add_nodes(self, num, data=None):
        raise DGLError("Readonly graph. Mutation is not allowed.")

    def add_edge(self, u, v, data=None):
        raise DGLError("Readonly graph. Mutation is not allowed.")

    def add_edges(self, u, v, data=None):
        raise DGLError("Readonly graph. Mutation is not allowed.")

    def segment_adjacency(self, typestr):
        """Membership Adjacency of the batch (cached): rows = graphs, columns = node (edge) ids."""
        if typestr not in self._segments:
            counts = self._batch_num_nodes if typestr == "nodes" else self._batch_num_edges
            self._segments[typestr] = segment_adjacency(counts, self._need_device())
        return self._segments[typestr]


def segment_adjacency(counts, device):
    """Adjacency whose row g holds the ids [sum(counts[:g]), sum(counts[:g+1])): forward = per-segment
    reduce, transposed = broadcast of a segment row to its members (the gradient)."""
    n_seg = len(counts)
    cnt = torch.tensor(list(counts), dtype=torch.int64)
    total = int(cnt.sum())
    if total >= (1 << 31):
        raise DGLError("segment_adjacency: more than 2^31 items")
    indptr = torch.zeros(n_seg + 1, dtype=torch.int64)
    indptr[1:] = torch.cumsum(cnt, 0)
    seg_of_item = torch.repeat_interleave(torch.arange(n_seg, dtype=torch.int32), cnt)
    csr_in = ops.CSR(n_seg, total, indptr.to(device), torch.arange(total, dtype=torch.int32, device=device), None)
    csr_out = ops.CSR(total, n_seg, torch.arange(total + 1, dtype=torch.int64, device=device),
                      seg_of_item.to(device), None)
    return Adjacency(n_seg, total, csr_in=csr_in, csr_out=csr_out)


def batch(graph_list, node_attrs=ALL, edge_attrs=ALL):
    """python/dgl/batched_graph.py:328-360."""
    return BatchedDGLGraph(graph_list, node_attrs, edge_attrs)


def unbatch(graph):
    """python/dgl/batched_graph.py:281-326: the component graphs with their slices of every field."""
    assert isinstance(graph, BatchedDGLGraph)
    src, dst = graph.edges() if graph.number_of_edges() > 0 else (None, None)
    out, n_off, e_off = [], 0, 0
    for n_i, e_i in zip(graph.batch_num_nodes, graph.batch_num_edges):
        g = DGLGraph()
        g.add_nodes(n_i)
        if e_i > 0:
            g.add_edges(src[e_off:e_off + e_i] - n_off, dst[e_off:e_off + e_i] - n_off)
        if graph._device is not None:
            g.to(graph._device)
        for k, v in graph.ndata.items():
            g.ndata[k] = v[n_off:n_off + n_i]
        for k, v in graph.edata.items():
            g.edata[k] = v[e_off:e_off + e_i]
        out.append(g)
        n_off += n_i
        e_off += e_i
    return out


# ---- readout ----------------------------------------------------------------------------------------

def _data(graph, typestr):
    return graph.ndata if typestr == "nodes" else graph.edata


def _segment_reduce(graph, typestr, feat, weight, red):
    """(n_graphs, ...) reduce of feat (optionally * weight) per graph: kernel (1) on the membership CSR."""
    adj = graph.segment_adjacency(typestr)
    flat = feat.reshape(feat.shape[0], -1)
    if weight is None:
        y = aggregate(adj, ops.MSG_COPY_SRC, red, x=flat)
    else:
        y = aggregate(adj, ops.MSG_SRC_MUL_EDGE, red, x=flat, w=weight.reshape(-1, 1).to(flat.dtype), w_width=1)
    return y.reshape((adj.n_dst,) + tuple(feat.shape[1:]))


def _sum_on(graph, typestr, feat, weight):
    """batched_graph.py:362-399."""
    data = _data(graph, typestr)
    f = data[feat]
    w = data[weight] if weight is not None else None
    if isinstance(graph, BatchedDGLGraph):
        return _segment_reduce(graph, typestr, f, w, ops.RED_SUM)
    if w is not None:
        f = w.reshape((-1,) + (1,) * (f.dim() - 1)) * f
    return f.sum(0)


def _mean_on(graph, typestr, feat, weight):
    """batched_graph.py:542-588: sum(w * feat) / sum(w), or the plain mean."""
    data = _data(graph, typestr)
    f = data[feat]
    w = data[weight] if weight is not None else None
    if isinstance(graph, BatchedDGLGraph):
        y = _segment_reduce(graph, typestr, f, w, ops.RED_SUM)
        if w is not None:
            den = _segment_reduce(graph, typestr, w.reshape(-1, 1).to(f.dtype), None, ops.RED_SUM)
        else:
            counts = graph.batch_num_nodes if typestr == "nodes" else graph.batch_num_edges
            den = torch.tensor(counts, dtype=f.dtype, device=f.device).reshape(-1, 1)
        return y / den.reshape((-1,) + (1,) * (f.dim() - 1))
    if w is None:
        return f.mean(0)
    return (w.reshape((-1,) + (1,) * (f.dim() - 1)) * f).sum(0) / w.sum(0)


def _max_on(graph, typestr, feat):
    """batched_graph.py:730-766: per-graph max; a graph without items reads 0."""
    f = _data(graph, typestr)[feat]
    if isinstance(graph, BatchedDGLGraph):
        return _segment_reduce(graph, typestr, f, None, ops.RED_MAX)
    return f.max(0)[0]


def sum_nodes(graph, feat, weight=None):
    """python/dgl/batched_graph.py:401-468."""
    return _sum_on(graph, "nodes", feat, weight)


def sum_edges(graph, feat, weight=None):
    """python/dgl/batched_graph.py:470-540."""
    return _sum_on(graph, "edges", feat, weight)


def mean_nodes(graph, feat, weight=None):
    """python/dgl/batched_graph.py:590-657."""
    return _mean_on(graph, "nodes", feat, weight)


def mean_edges(graph, feat, weight=None):
    """python/dgl/batched_graph.py:659-728."""
    return _mean_on(graph, "edges", feat, weight)


def max_nodes(graph, feat):
    """python/dgl/batched_graph.py:768-792."""
    return _max_on(graph, "nodes", feat)


def max_edges(graph, feat):
    """python/dgl/batched_graph.py:794-815."""
    return _max_on(graph, "edges", feat)
