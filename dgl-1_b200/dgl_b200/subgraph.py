"""DGLSubGraph: a read-only graph that remembers where its nodes and edges came from
(python/dgl/subgraph.py:9-146).  Produced by the neighbour sampler (contrib/sampling)."""
import torch

from .base import DGLError
from .graph import DGLGraph, _as_ids


class DGLSubGraph(DGLGraph):
    def __init__(self, parent, parent_nid, parent_eid, src, dst, layer_ids=None, shared=False):
        """parent_nid / parent_eid: int64 device tensors mapping subgraph ids to parent ids; (src, dst):
        the subgraph's edges in subgraph-edge-id order, compacted ids."""
        super(DGLSubGraph, self).__init__()
        if shared:
            raise DGLError("Shared mode is not yet supported.")          # subgraph.py:48-49
        self._parent = parent
        self._parent_nid = parent_nid
        self._parent_eid = parent_eid
        self._layer_ids = layer_ids
        self._device = parent_nid.device if parent_nid.is_cuda else parent._device
        DGLGraph.add_nodes(self, int(parent_nid.shape[0]))
        if src.numel():
            DGLGraph.add_edges(self, src, dst)
        self._readonly = True
        self._frozen = True

    @property
    def parent_nid(self):
        """Subgraph node id -> parent node id (subgraph.py:66-78)."""
        return self._parent_nid

    @property
    def parent_eid(self):
        """Subgraph edge id -> parent edge id (subgraph.py:89-101)."""
        return self._parent_eid

    @property
    def layer_ids(self):
        """Hop at which every node was reached (0 = seed): SampledSubgraph::layer_ids
        (src/graph/immutable_graph.cc:880-895)."""
        return self._layer_ids

    def copy_from_parent(self):
        """Node / edge features of the parent, rows of this subgraph (subgraph.py:117-127).  Old features
        are dropped."""
        self.ndata.clear()
        self.edata.clear()
        for k, v in self._parent.ndata.items():
            self.ndata[k] = v[self._parent_nid.to(v.device)]
        for k, v in self._parent.edata.items():
            self.edata[k] = v[self._parent_eid.to(v.device)]

    def copy_to_parent(self, inplace=False):
        """Write this subgraph's features back into the parent's rows (subgraph.py:103-115)."""
        for feats, pfeats, idx in ((self.ndata, self._parent.ndata, self._parent_nid),
                                   (self.edata, self._parent.edata, self._parent_eid)):
            for k, v in feats.items():
                if k not in pfeats:
                    base = torch.zeros((pfeats.num_rows(),) + tuple(v.shape[1:]), dtype=v.dtype, device=v.device)
                else:
                    base = pfeats[k]
                i = idx.to(v.device)
                if inplace:
                    base[i] = v
                    dict.__setitem__(pfeats, k, base)
                else:
                    dict.__setitem__(pfeats, k, base.index_put((i,), v))

    def map_to_subgraph_nid(self, parent_vids):
        """Parent node ids -> ids in this subgraph, -1 where the node is not part of it
        (GraphOp::MapParentIdToSubgraphId, src/graph/graph_op.cc; the induced ids are sorted)."""
        v = _as_ids(parent_vids, self._parent_nid.device)
        nid = self._parent_nid
        if nid.numel() == 0:
            return torch.full_like(v, -1)
        pos = torch.searchsorted(nid, v).clamp_(max=nid.numel() - 1)
        return torch.where(nid[pos] == v, pos, torch.full_like(pos, -1))
