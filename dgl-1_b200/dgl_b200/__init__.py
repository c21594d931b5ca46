"""dgl_b200: B200-native engine for DGL 0.1.x's builtin message-passing path.

Host side mirrors the reference's interface for that path (python/dgl/graph.py update_all /
send_and_recv / pull, python/dgl/function); all computation is in libdgl_b200.so
(include/dgl_b200.h).  Importing this package does not need a GPU; calling an op does.
"""
from .base import ALL, DGLError
from . import function
from . import init
from . import ops
from .graph import DGLGraph
from .udf import EdgeBatch, NodeBatch
from .batched_graph import (BatchedDGLGraph, batch, unbatch, sum_nodes, sum_edges, mean_nodes,
                            mean_edges, max_nodes, max_edges)
from .edge_ops import sddmm, edge_softmax, gat_attention
from .propagate import prop_nodes, prop_edges, prop_nodes_bfs, prop_nodes_topo
from .subgraph import DGLSubGraph
from . import contrib
from .traversal import bfs_nodes_generator, bfs_edges_generator, topological_nodes_generator

__version__ = "0.1.0"
