"""Per-kernel breakdown of one forward + backward of the differentiable GAT edge path at C2 size (8 heads x 16), slot
order: torch.profiler over one step (CUDA activities).  python scripts/gat_chain_breakdown.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "dgl-1_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

from benchmarks import graphs  # noqa: E402
from dgl_b200 import autograd, edge_ops  # noqa: E402

DEV = torch.device("cuda")
c = graphs.C2_REDDIT
n, e, heads, hid = c["n"], c["e"], 8, 16
src, dst = graphs.powerlaw_symmetric(n, e, c["seed"], DEV)
adj = autograd.Adjacency(n, n, coo=(dst, src, None))
_ = adj.csr_in, adj.csr_out
adj.slot_view()
adj.drop_coo()
del src, dst
a1 = torch.randn(n, heads, device=DEV, requires_grad=True)
a2 = torch.randn(n, heads, device=DEV, requires_grad=True)
ft = torch.randn(n, heads, hid, device=DEV, requires_grad=True)
gout = torch.randn(n, heads, hid, device=DEV)


def step():
    out, _ = edge_ops.gat_attention(adj, a1, a2, ft, 0.2, want_alpha=False, slot_order=True)
    out.backward(gout)
    a1.grad = a2.grad = ft.grad = None


for _ in range(2):
    step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    step()
    torch.cuda.synchronize()
rows = sorted(prof.key_averages(), key=lambda k: -k.device_time_total)[:16]
tot = sum(k.device_time_total for k in prof.key_averages())
print("total device time %.2f ms" % (tot / 1e3))
for k in rows:
    print("%-90s n=%d %8.2f ms" % (k.key[:90], k.count, k.device_time_total / 1e3))
