"""Generate tests/golden/sampling_*.npz by running the UNMODIFIED reference's neighbour sampler
(ImmutableGraph::NeighborUniformSample, src/graph/immutable_graph.cc:998-1008, through
graph_index._CAPI_DGLGraphUniformSampling) on its own read-only graph.  TEST INFRASTRUCTURE ONLY; build
container only:

    DGL_LIBRARY_PATH=oracle/_ref PYTHONPATH=/root/reference/python python oracle/gen_golden_sampling.py

The reference seeds rand_r with time(nullptr) (:778): only calls whose fanout covers every degree they meet
are deterministic ("exact": 1 - compared bit for bit).  The other records ("exact": 0) are REAL outputs of the
reference kept to pin oracle.check_neighbor_subgraph, the property checker the CUDA sampler is held to.

Per case: the graph (src, dst in edge-id order, n), seeds, fanout, hops, neighbour type; induced_nodes,
layer_ids, induced_edges, and the subgraph's edges in subgraph-edge-id order (sub_src, sub_dst).
"""
import os

import numpy as np
import scipy.sparse as sp
import torch

import dgl
from dgl import graph_index as gi
from dgl import utils

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")


def sample(g, seeds, fanout, hops, ntype):
    ids = [utils.toindex(torch.from_numpy(np.asarray(seeds, dtype=np.int64))).todgltensor()]
    rst = gi._uniform_sampling(g._graph, ids, ntype, hops, fanout)
    sub = gi.SubgraphIndex(rst(0), g._graph, utils.toindex(rst(1)), utils.toindex(rst(2)))
    layers = utils.toindex(rst(3)).tonumpy()
    s, d, e = sub.edges()
    s, d, e = s.tonumpy(), d.tonumpy(), e.tonumpy()
    order = np.argsort(e, kind="stable")
    assert np.array_equal(e[order], np.arange(len(e)))
    return dict(induced_nodes=sub.induced_nodes.tonumpy().astype(np.int64),
                induced_edges=sub.induced_edges.tonumpy().astype(np.int64),
                layer_ids=layers.astype(np.int64), sub_src=s[order].astype(np.int64),
                sub_dst=d[order].astype(np.int64))


def main():
    idx = 0
    for gseed, (n, density) in enumerate([(100, 0.1), (300, 0.04)]):
        arr = (sp.random(n, n, density=density, format="coo", random_state=gseed + 11) != 0).astype(np.int64)
        g = dgl.DGLGraph(arr, readonly=True)
        src, dst, eid = g.edges(form="all")
        src, dst, eid = src.numpy(), dst.numpy(), eid.numpy()
        order = np.argsort(eid, kind="stable")
        src, dst = src[order], dst[order]                      # edge-id order
        max_deg = int(max(np.bincount(src, minlength=n).max(), np.bincount(dst, minlength=n).max()))
        rng = np.random.RandomState(5 + gseed)
        cases = []
        for ntype in ("in", "out"):
            cases += [([int(rng.randint(n))], max_deg, 1, ntype, 1),
                      ([int(rng.randint(n))], max_deg + 50, 2, ntype, 1),
                      (rng.choice(n, 10, replace=False).tolist(), max_deg, 2, ntype, 1),
                      (rng.randint(0, n, 12).tolist() + [3, 3], max_deg, 3, ntype, 1),   # duplicate seeds
                      (rng.choice(n, 10, replace=False).tolist(), 5, 2, ntype, 0),
                      (rng.choice(n, 6, replace=False).tolist(), max(max_deg - 3, 1), 2, ntype, 0),  # drop-set branch
                      (rng.choice(n, 4, replace=False).tolist(), 1, 3, ntype, 0),
                      (list(range(n)), 3, 1, ntype, 0)]
        for seeds, fanout, hops, ntype, exact in cases:
            rec = sample(g, seeds, fanout, hops, ntype)
            np.savez_compressed(os.path.join(OUT, "sampling_%d.npz" % idx), n=n, src=src, dst=dst,
                                seeds=np.asarray(seeds, dtype=np.int64), fanout=fanout, hops=hops,
                                ntype=ntype, exact=exact, **rec)
            idx += 1
    print("wrote %d sampling cases" % idx)


if __name__ == "__main__":
    main()
