"""Shared helpers for the tests: seeded synthetic graphs in the shapes the reference's own
tests use (tests/compute/test_specialization.py:9-21,361-370,517-530)."""
import numpy as np


def random_graph(n, e, seed, hubs=0, hub_deg=0, multigraph=True, zero_deg_frac=0.0):
    """Directed edge list (src, dst) int64 with optional hub destinations (power-law stand-in),
    parallel edges (multigraph) and a fraction of nodes that never appear as destination."""
    rng = np.random.RandomState(seed)
    n_dst = max(1, int(n * (1.0 - zero_deg_frac)))
    src = rng.randint(0, n, size=e).astype(np.int64)
    dst = rng.randint(0, n_dst, size=e).astype(np.int64)
    if hubs:
        hs = rng.randint(0, n, size=hubs * hub_deg).astype(np.int64)
        hd = np.repeat(rng.choice(n_dst, size=hubs, replace=False), hub_deg).astype(np.int64)
        src = np.concatenate([src, hs])
        dst = np.concatenate([dst, hd])
        perm = rng.permutation(len(src))
        src, dst = src[perm], dst[perm]
    if not multigraph:
        key = np.unique(src * n + dst)
        rng.shuffle(key)
        src, dst = key // n, key % n
    return src, dst


def hub_graph():
    """The 10-node / 17-edge graph of the reference's tests/compute/test_function.py:5-19:
    0 -> 1..8 -> 9, plus 9 -> 0."""
    src, dst = [], []
    for i in range(1, 9):
        src += [0, i]
        dst += [i, 9]
    src.append(9)
    dst.append(0)
    return np.asarray(src, dtype=np.int64), np.asarray(dst, dtype=np.int64)


# edge feature of that graph, tests/compute/test_function.py:16-17
HUB_EDGE_H = [1., 2., 1., 3., 1., 4., 1., 5., 1., 6., 1., 7., 1., 8., 1., 9., 10.]
