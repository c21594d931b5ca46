"""Neighbour sampling on the device: ImmutableGraph::NeighborUniformSample
(src/graph/immutable_graph.cc:772-1008) level by level instead of vertex by vertex.

The reference pops one vertex at a time from a BFS queue, draws its neighbours, and looks every neighbour up
in a hash set.  Layers of that queue are independent of the order inside a layer, so here a layer is ONE
launch of the sampling kernel (csrc/sampling.cu: a warp per frontier vertex) which also does the "seen before?"
test - an atomicCAS on a per-vertex layer array instead of a hash set - and appends the next layer's frontier.
The discovered set is kept as a bitmap; its popcount scan gives both the ascending vertex list and the
compaction map, so no sort is needed anywhere.  Outputs have the reference's shape exactly:
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


def sample_subgraph(csr, seeds, fanout, num_hops, seed=0, neighbor_type="in", state=None):
    """`csr`: the (row, column)-sorted CSR of the parent (DGLGraph.immutable_csr); `seeds` int64 ids on its
    device (duplicates allowed, :787-793); `fanout` = the reference's expand_factor / num_neighbor.
    `state`: an ops.SampleState of the parent to reuse across calls (one is made, and cached on the CSR,
    otherwise).  The whole call is dglb200_sample_subgraph: per hop the offsets of the frontier's picks (one
    sync: their number sizes the outputs) and the sampling kernel, which also discovers the next frontier; at
    the end the discovered set is compacted out of the bitmap and the picks are placed into the subgraph's CSR
    (one more sync: the vertex count)."""
    if int(fanout) < 0 or int(num_hops) < 0:
        raise DGLError("sample_subgraph: fanout and num_hops must be >= 0")
    dev = csr.device
    n = csr.n_rows
    seeds = seeds.to(dev).to(torch.int64).reshape(-1).contiguous()
    if seeds.numel() == 0:
        raise DGLError("sample_subgraph: no seed vertices")
    if state is None:
        state = getattr(csr, "_sample_state", None)
        if state is None:
            state = csr._sample_state = ops.SampleState(n, dev)
    nodes, layers, sub_indptr, sub_indices, sub_rows, induced = ops.sample_subgraph(csr, state, seeds, fanout,
                                                                                    int(num_hops), seed)
    n_sub = int(nodes.shape[0])
    sub = ops.CSR(n_sub, n_sub, sub_indptr, sub_indices,
                  torch.arange(sub_indices.shape[0], dtype=torch.int32, device=dev))     # :948-949
    return SampledSubgraph(nodes, layers, induced, sub, sub_rows, neighbor_type)
