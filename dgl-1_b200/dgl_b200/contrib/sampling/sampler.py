"""NeighborSampler: same arguments and iteration protocol as the reference's
(python/dgl/contrib/sampling/sampler.py:18-82, 196-259), the sampling itself on the device.

The reference hands `num_workers` seed batches per call to _CAPI_DGLGraphUniformSampling*, which samples them
on OpenMP threads (src/graph/graph_apis.cc:434-457); here a batch is a handful of kernel launches on the
current stream, so `num_workers` only sets how many batches are sampled ahead, and `prefetch` needs no
thread."""
import torch

from ... import sampling as engine
from ...base import DGLError
from ...graph import _as_ids
from ...subgraph import DGLSubGraph

__all__ = ["NeighborSampler"]


class NSSubgraphLoader(object):
    def __init__(self, g, batch_size, expand_factor, num_hops=1, neighbor_type="in", node_prob=None,
                 seed_nodes=None, shuffle=False, num_workers=1, return_seed_id=False, seed=None):
        self._g = g
        if not g.is_readonly():
            raise NotImplementedError("subgraph loader only support read-only graphs.")   # sampler.py:24-25
        if node_prob is not None:
            # unreachable in the reference too: graph_index.py:674 calls `_nonuniform_sampling`, defined nowhere, and no
            # PackedFunc is registered for GetNonUniformSample (immutable_graph.cc:746-770; graph_apis.cc:434-472)
            raise NotImplementedError("node_prob: only uniform neighbour sampling runs on the device")
        if neighbor_type not in ("in", "out"):
            raise DGLError("neighbor_type must be 'in' or 'out'")
        if not isinstance(expand_factor, int):
            raise NotImplementedError("expand_factor: only an integer neighbour count is supported "
                                      "(like the reference's C API, graph_apis.cc:444)")
        self._batch_size = int(batch_size)
        self._expand_factor = expand_factor
        self._num_hops = int(num_hops)
        self._return_seed_id = return_seed_id
        dev = g._need_device()
        if seed_nodes is None:
            self._seed_nodes = torch.arange(g.number_of_nodes(), dtype=torch.int64, device=dev)
        else:
            self._seed_nodes = _as_ids(seed_nodes, dev)
        if shuffle:
            self._seed_nodes = self._seed_nodes[torch.randperm(self._seed_nodes.shape[0], device=dev)]
        self._num_workers = int(num_workers)
        self._neighbor_type = neighbor_type
        self._subgraphs = []
        self._seed_ids = []
        self._subgraph_idx = 0
        # the reference seeds rand_r with time() per call (immutable_graph.cc:778); a fixed `seed` makes an
        # epoch reproducible, None draws one from torch's generator
        self._seed = int(torch.randint(0, 2 ** 62, (1,)).item()) if seed is None else int(seed)

    def _prefetch(self):
        csr = self._g.immutable_csr(self._neighbor_type)
        num_nodes = int(self._seed_nodes.shape[0])
        for _ in range(self._num_workers):
            start = self._subgraph_idx * self._batch_size
            if start >= num_nodes:
                break
            end = min(start + self._batch_size, num_nodes)
            seeds = self._seed_nodes[start:end]
            sub = engine.sample_subgraph(csr, seeds, self._expand_factor, self._num_hops,
                                         seed=self._seed + self._subgraph_idx, neighbor_type=self._neighbor_type)
            src, dst = sub.edges()
            self._subgraphs.append(DGLSubGraph(self._g, sub.induced_nodes, sub.induced_edges, src, dst,
                                               layer_ids=sub.layer_ids))
            if self._return_seed_id:
                self._seed_ids.append(seeds)
            self._subgraph_idx += 1

    def __iter__(self):
        return self

    def __next__(self):
        if len(self._subgraphs) == 0:
            self._prefetch()
        if len(self._subgraphs) == 0:
            raise StopIteration
        aux_infos = {}
        if self._return_seed_id:
            aux_infos["seeds"] = self._seed_ids.pop(0)
        return self._subgraphs.pop(0), aux_infos

    next = __next__


def NeighborSampler(g, batch_size, expand_factor, num_hops=1, neighbor_type="in", node_prob=None,
                    seed_nodes=None, shuffle=False, num_workers=1, return_seed_id=False, prefetch=False,
                    seed=None):
    """Iterator of (DGLSubGraph, aux) over batches of `batch_size` seed nodes, each the `num_hops`-hop
    neighbourhood with at most `expand_factor` sampled neighbours per vertex (sampler.py:196-259).
    `prefetch` is accepted and ignored: launches are asynchronous already."""
    return NSSubgraphLoader(g, batch_size, expand_factor, num_hops, neighbor_type, node_prob, seed_nodes,
                            shuffle, num_workers, return_seed_id, seed)
