"""Frontier generators for message propagation, computed level-synchronously on the device.

Mirrors python/dgl/traversal.py (bfs_nodes_generator :12-49, bfs_edges_generator :51-89,
topological_nodes_generator :91-126) whose C++ is src/graph/traversal.h:26-212 (a sequential queue
walk).  Frontiers are returned in the reference's order: a node (edge) enters a frontier at the
position at which the queue walk discovers it - the parents of a frontier are visited in frontier
order and their out-edges in ascending edge id, so "discovery order" is the first (BFS) or last
(topological: the decrement that reaches zero) occurrence in the concatenated out-edge lists that
DGLGraph.out_edges returns.  Integer id plumbing only; the per-level unique/relabel is kernel (4b)'s
dglb200_unique_sorted.
"""
import torch

from . import ops
from .base import DGLError
from .graph import _as_ids


def _next_hops(graph, frontier, reverse):
    """(neighbour, edge id) lists of the frontier: successors along out-edges, or predecessors along
    in-edges when reverse (Graph::SuccVec / PredVec, OutEdgeVec / InEdgeVec)."""
    if reverse:
        u, _, eid = graph.in_edges(frontier)
        return u, eid
    _, v, eid = graph.out_edges(frontier)
    return v, eid


def _order_by(uniq, key):
    return uniq[torch.sort(key, stable=True)[1]]


def bfs_nodes_generator(graph, source, reverse=False):
    """Node frontiers of a breadth-first search from `source` (traversal.h:26-60)."""
    dev = graph._need_device()
    n = graph.number_of_nodes()
    frontier = _as_ids(source, dev)
    visited = torch.zeros(n, dtype=torch.bool, device=dev)
    visited[frontier] = True
    frontiers = [frontier]
    while True:
        nxt, _ = _next_hops(graph, frontier, reverse)
        nxt = nxt[~visited[nxt]]
        if nxt.numel() == 0:
            return frontiers
        uniq, inv = ops.unique_sorted(nxt, n)
        pos = torch.arange(nxt.numel(), dtype=torch.int64, device=dev)
        first = torch.full((uniq.numel(),), nxt.numel(), dtype=torch.int64, device=dev)
        first.scatter_reduce_(0, inv, pos, "amin")
        frontier = _order_by(uniq, first)
        visited[frontier] = True
        frontiers.append(frontier)


def bfs_edges_generator(graph, source, reverse=False):
    """Edge frontiers of a breadth-first search: the tree edge through which every node is first
    reached (traversal.h:78-118)."""
    dev = graph._need_device()
    n = graph.number_of_nodes()
    frontier = _as_ids(source, dev)
    visited = torch.zeros(n, dtype=torch.bool, device=dev)
    visited[frontier] = True
    frontiers = []
    while True:
        nxt, eid = _next_hops(graph, frontier, reverse)
        new = ~visited[nxt]
        nxt, eid = nxt[new], eid[new]
        if nxt.numel() == 0:
            return frontiers
        uniq, inv = ops.unique_sorted(nxt, n)
        pos = torch.arange(nxt.numel(), dtype=torch.int64, device=dev)
        first = torch.full((uniq.numel(),), nxt.numel(), dtype=torch.int64, device=dev)
        first.scatter_reduce_(0, inv, pos, "amin")
        first = torch.sort(first)[0]
        frontier = nxt[first]
        visited[frontier] = True
        frontiers.append(eid[first])


def topological_nodes_generator(graph, reverse=False):
    """Node frontiers of Kahn's algorithm (traversal.h:174-212); raises on a cycle."""
    dev = graph._need_device()
    n = graph.number_of_nodes()
    deg = (graph.out_degrees() if reverse else graph.in_degrees()).clone()
    frontier = (deg == 0).nonzero().reshape(-1)
    visited = int(frontier.numel())
    frontiers = [frontier]
    while frontier.numel() > 0:
        nxt, _ = _next_hops(graph, frontier, reverse)
        if nxt.numel() == 0:
            break
        uniq, inv = ops.unique_sorted(nxt, n)
        deg[uniq] -= torch.bincount(inv, minlength=uniq.numel())
        done = deg[uniq] == 0
        pos = torch.arange(nxt.numel(), dtype=torch.int64, device=dev)
        last = torch.full((uniq.numel(),), -1, dtype=torch.int64, device=dev)
        last.scatter_reduce_(0, inv, pos, "amax")
        frontier = _order_by(uniq[done], last[done])
        if frontier.numel() == 0:
            break
        visited += int(frontier.numel())
        frontiers.append(frontier)
    if visited != n:
        raise DGLError("Error in topological traversal: loop detected in the given graph.")
    return frontiers
