"""Neighbour sampling on the device: ImmutableGraph::NeighborUniformSample
(src/graph/immutable_graph.cc:772-1008) level by level instead of vertex by vertex.

The reference pops one vertex at a time from a BFS queue, draws its neighbours, and looks every neighbour up
in a hash set.  Layers of that queue are independent of the order inside a layer, so here a layer is ONE
launch of the sampling kernel (csrc/sampling.cu: a warp per frontier vertex) and the "seen before?" test is a
sorted-set union (dglb200_unique_sorted, kernel 4).  Outputs have the reference's shape exactly:
induced vertices ascending, their layer ids, the parent edge id of every subgraph edge, and the subgraph's
CSR with compacted column ids and edge ids 0..E-1 (:875-955, CompactSubgraph :963-983).
"""
import torch

from . import ops
from .base import DGLError


class SampledSubgraph(object):
    """What SampleSubgraph returns (include/dgl/graph_interface.h SampledSubgraph): all int64, on the
    device.  `csr` is keyed by the vertex whose neighbour list was sampled (the destination for
    neighbor_type "in"); sub_rows / sub_cols are its edges expanded, in subgraph-edge-id order."""

    def __init__(self, induced_nodes, layer_ids, induced_edges, csr, sub_rows, neighbor_type):
        self.induced_nodes = induced_nodes
        self.layer_ids = layer_ids
        self.induced_edges = induced_edges
        self.csr = csr
        self.sub_rows = sub_rows
        self.sub_cols = csr.indices.long()
        self.neighbor_type = neighbor_type

    def edges(self):
        """(src, dst) of the subgraph's edges in subgraph-edge-id order, compacted ids."""
        if self.neighbor_type == "in":
            return self.sub_cols, self.sub_rows
        return self.sub_rows, self.sub_cols


def sample_subgraph(csr, seeds, fanout, num_hops, seed=0, neighbor_type="in"):
    """`csr`: the (row, column)-sorted CSR of the parent (DGLGraph.immutable_csr); `seeds` int64 ids on its
    device (duplicates allowed, :787-793); `fanout` = the reference's expand_factor / num_neighbor."""
    if int(fanout) < 0 or int(num_hops) < 0:
        raise DGLError("sample_subgraph: fanout and num_hops must be >= 0")
    dev = csr.device
    n = csr.n_rows
    seeds = seeds.to(dev).to(torch.int64).reshape(-1).contiguous()
    if seeds.numel() == 0:
        raise DGLError("sample_subgraph: no seed vertices")
    if int(seeds.min()) < 0 or int(seeds.max()) >= n:
        raise DGLError("Invalid vertex: a seed id is outside [0, %d)" % n)
    nodes, _ = ops.unique_sorted(seeds, n, want_inverse=False)         # layer 0, ascending
    layers = torch.zeros(nodes.shape[0], dtype=torch.int64, device=dev)
    frontier = nodes
    rows_l, cols_l, eids_l = [], [], []
    for hop in range(int(num_hops)):
        if frontier.numel() == 0:
            break
        _, r, c, e = ops.neighbor_sample(csr, frontier, fanout, seed, want=("row", "col", "eid"))
        rows_l.append(r)
        cols_l.append(c)
        eids_l.append(e)
        if c.numel() == 0:
            break
        # union of the known vertices and this layer's picks; what is new forms the next frontier
        known = nodes.shape[0]
        merged, inv = ops.unique_sorted(torch.cat([nodes, c]), n)
        old_pos = inv[:known]
        new_layers = torch.full((merged.shape[0],), hop + 1, dtype=torch.int64, device=dev)
        new_layers[old_pos] = layers
        is_new = torch.ones(merged.shape[0], dtype=torch.bool, device=dev)
        is_new[old_pos] = False
        frontier = merged[is_new]
        nodes, layers = merged, new_layers
    n_sub = int(nodes.shape[0])
    if rows_l:
        rows = torch.cat(rows_l)
        cols = torch.cat(cols_l)
        eids = torch.cat(eids_l)
    else:
        rows = cols = eids = torch.empty(0, dtype=torch.int64, device=dev)
    # compaction (nodes is sorted and holds every id) + the subgraph's CSR: rows ascending, the picks of a
    # row in the order the kernel emitted them (ascending slot) - a stable build keyed by the row
    sub_r = torch.searchsorted(nodes, rows)
    sub_c = torch.searchsorted(nodes, cols)
    sub = ops.csr_build(sub_r, sub_c, n_sub, n_sub, order=0)
    perm = sub.eids.long()
    induced_edges = eids[perm]
    sub.eids = torch.arange(sub.nnz, dtype=torch.int32, device=dev)   # :948-949
    sub_rows = ops.csr_row_ids(sub) if sub.nnz else torch.empty(0, dtype=torch.int64, device=dev)
    return SampledSubgraph(nodes, layers, induced_edges, sub, sub_rows, neighbor_type)
