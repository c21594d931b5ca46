"""Deterministic synthetic graphs in the shapes BASELINE.json names.
Bench / test infrastructure (no dataset download is possible); edge lists are generated with
torch on the requested device and returned as int64 (src, dst) in edge-id order."""
import math

import torch

C1_CORA = dict(n=2708, e=10556, feat_in=1433, hidden=16, classes=7, seed=0)
C2_REDDIT = dict(n=232965, e=114615892, feat_in=602, hidden=256, classes=41, seed=1)
C3_PUBMED = dict(n=19717, e=88648, feat_in=500, heads=8, hidden=8, seed=2)
C4_AIFB = dict(n=8285, e=29043, rels=90, hidden=16, seed=3)
C5_RMAT = dict(n=50_000_000, e=150_000_000, feat=128, seed=4)


def _gen(device, seed):
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    return g


def _powerlaw_cdf(n, alpha, max_p, device):
    """CDF of p_i ~ (i + i0)^-alpha with i0 chosen (bisection) so that p_0 == max_p."""
    i = torch.arange(n, dtype=torch.float64, device=device)
    lo, hi = 0.0, float(n) * 100
    for _ in range(80):
        mid = 0.5 * (lo + hi)
        w = (i + 1.0 + mid).pow(-alpha)
        p0 = float(w[0] / w.sum())
        if p0 > max_p:
            lo = mid
        else:
            hi = mid
    w = (i + 1.0 + hi).pow(-alpha)
    return torch.cumsum(w / w.sum(), 0)


def powerlaw_symmetric(n, e, seed, device, alpha=0.8, max_deg=22000):
    """Reddit-shaped graph: e/2 pairs (a, b) with a ~ truncated power law, b ~ uniform, then
    both directions are kept (symmetrised), so in- and out-degrees share one heavy-tailed
    sequence with maximum ~max_deg and mean e/n.  Node labels are randomly permuted."""
    half = e // 2
    g = _gen(device, seed)
    cdf = _powerlaw_cdf(n, alpha, max_deg / float(half) - 1.0 / n, device).float()
    perm = torch.randperm(n, generator=g, device=device)
    a = torch.empty(half, dtype=torch.int64, device=device)
    b = torch.empty(half, dtype=torch.int64, device=device)
    step = 1 << 24
    for s in range(0, half, step):
        m = min(step, half - s)
        u = torch.rand(m, generator=g, device=device)
        a[s:s + m] = perm[torch.searchsorted(cdf, u).clamp_(max=n - 1)]
        b[s:s + m] = torch.randint(0, n, (m,), generator=g, device=device)
    src = torch.cat([a, b])
    dst = torch.cat([b, a])
    if e % 2:
        src = torch.cat([src, src[:1]])
        dst = torch.cat([dst, dst[:1]])
    return src, dst


def uniform_graph(n, e, seed, device, self_loops=False):
    """Erdos-Renyi-like directed multigraph (Cora / PubMed / AIFB stand-ins) (+ n self loops as
    examples/pytorch/gcn/gcn_spmv.py:135-136 and gat/train.py:186-189 add)."""
    g = _gen(device, seed)
    src = torch.randint(0, n, (e,), generator=g, device=device)
    dst = torch.randint(0, n, (e,), generator=g, device=device)
    if self_loops:
        loop = torch.arange(n, device=device)
        src = torch.cat([src, loop])
        dst = torch.cat([dst, loop])
    return src, dst


def rmat(scale, e, seed, device, n=None, abcd=(0.57, 0.19, 0.19, 0.05)):
    """R-MAT edge list (no dedup) on 2^scale vertices, truncated to n by modulo, labels randomly
    permuted."""
    g = _gen(device, seed)
    a, b, c, _ = abcd
    src = torch.zeros(e, dtype=torch.int64, device=device)
    dst = torch.zeros(e, dtype=torch.int64, device=device)
    for _ in range(scale):
        u = torch.rand(e, generator=g, device=device)
        sbit = (u >= a + b).long()
        dbit = ((u >= a) & (u < a + b) | (u >= a + b + c)).long()
        src = src * 2 + sbit
        dst = dst * 2 + dbit
        del u, sbit, dbit
    n = n or (1 << scale)
    perm = torch.randperm(1 << scale, generator=g, device=device)
    src = perm[src] % n
    dst = perm[dst] % n
    return src, dst


def spmm_alg_bytes(e, n_out, feat, elem=4, idx=4, ptr=8, weights=False):
    """ALGORITHMIC bytes of one SpMM (SURVEY.md 8d gather model): every edge reads one
    feat-wide source row and its column id, every output row is written once, the row pointer
    is read once (int64 here, so ptr = 8)."""
    b = e * (feat * elem + idx) + n_out * feat * elem + (n_out + 1) * ptr
    if weights:
        b += e * (4 + idx)
    return b


def spmm_min_bytes(e, n_in, n_out, feat, elem=4, idx=4, ptr=8):
    """Compulsory DRAM bytes: every input / output row once + the structure."""
    return (n_in + n_out) * feat * elem + e * idx + (n_out + 1) * ptr
