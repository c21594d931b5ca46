"""Message propagation over frontiers (python/dgl/propagate.py:9-146): one pull (node frontiers)
or send_and_recv (edge frontiers) per level, each a device-side subset schedule."""
from . import traversal as trv

__all__ = ["prop_nodes", "prop_edges", "prop_nodes_bfs", "prop_nodes_topo"]


def prop_nodes(graph, nodes_generator, message_func="default", reduce_func="default",
               apply_node_func="default"):
    graph.prop_nodes(nodes_generator, message_func, reduce_func, apply_node_func)


def prop_edges(graph, edges_generator, message_func="default", reduce_func="default",
               apply_node_func="default"):
    graph.prop_edges(edges_generator, message_func, reduce_func, apply_node_func)


def prop_nodes_bfs(graph, source, reverse=False, message_func="default", reduce_func="default",
                   apply_node_func="default"):
    """python/dgl/propagate.py:57-86."""
    prop_nodes(graph, trv.bfs_nodes_generator(graph, source, reverse), message_func, reduce_func,
               apply_node_func)


def prop_nodes_topo(graph, reverse=False, message_func="default", reduce_func="default",
                    apply_node_func="default"):
    """python/dgl/propagate.py:88-114."""
    prop_nodes(graph, trv.topological_nodes_generator(graph, reverse), message_func, reduce_func,
               apply_node_func)
