"""-m gpu: the differentiable edge operators (dgl_b200/edge_ops.py: SDDMM u-op-v, edge softmax, GAT
attention) and the builtin edge functions of apply_edges, forward AND gradients, against the
reference's formulation of the same expressions: per-edge gathers + torch autograd, which is what
apply_edges + an edge UDF + update_all do in the reference (examples/pytorch/gat/train.py:61-97;
runtime/scheduler.py:234-277).  fp32, rtol 1e-4 (the sums are re-associated).

tests/test_host_layer_cpu.py runs the same bodies on CPU tensors with tests/cpu_engine.py behind
`ops` (autograd wiring only)."""
import numpy as np
import pytest
import torch

from tests.util import random_graph

pytestmark = pytest.mark.gpu
RTOL = 1e-4


def device():
    return torch.device("cuda")


def close(got, want, rtol=RTOL):
    got, want = got.detach().float().cpu(), want.detach().float().cpu()
    assert got.shape == want.shape, (got.shape, want.shape)
    tol = rtol * (float(want.abs().max()) + 1e-30)
    assert float((got - want).abs().max()) <= tol + rtol * float(want.abs().max()), float((got - want).abs().max())


def _setup(seed, n=70, e=500, zero_deg_frac=0.1):
    import dgl_b200
    src, dst = random_graph(n, e, seed, hubs=1, hub_deg=90, zero_deg_frac=zero_deg_frac)
    g = dgl_b200.DGLGraph()
    g.add_nodes(n)
    g.add_edges(torch.from_numpy(src), torch.from_numpy(dst))
    g.to(device())
    return g, torch.from_numpy(src).to(device()), torch.from_numpy(dst).to(device())


def _leaf(*shape, seed=0):
    gen = torch.Generator().manual_seed(seed)
    return torch.randn(*shape, generator=gen).to(device()).requires_grad_(True)


@pytest.mark.parametrize("op", ["u_add_v", "u_mul_v", "u_dot_v", "copy_u", "copy_v"])
@pytest.mark.parametrize("tail", [(6,), (4, 8), (3, 1)])
def test_sddmm_forward_and_gradients(op, tail):
    import dgl_b200
    g, src, dst = _setup(1)
    n, e = g.number_of_nodes(), g.number_of_edges()
    u, v = _leaf(n, *tail, seed=1), _leaf(n, *tail, seed=2)
    u2, v2 = u.detach().clone().requires_grad_(True), v.detach().clone().requires_grad_(True)
    if op == "u_add_v":
        want = u2[src] + v2[dst]
    elif op == "u_mul_v":
        want = u2[src] * v2[dst]
    elif op == "u_dot_v":
        want = (u2[src] * v2[dst]).sum(-1, keepdim=True)
    elif op == "copy_u":
        want = u2[src] + 0 * v2.sum()
    else:
        want = v2[dst] + 0 * u2.sum()
    got = dgl_b200.sddmm(g, op, None if op == "copy_v" else u, None if op == "copy_u" else v)
    close(got, want)
    gout = torch.randn(want.shape, generator=torch.Generator().manual_seed(3)).to(device())
    got.backward(gout)
    want.backward(gout)
    if op != "copy_v":
        close(u.grad, u2.grad)
    if op != "copy_u":
        close(v.grad, v2.grad)


def test_edge_softmax_forward_and_gradient():
    import dgl_b200
    g, src, dst = _setup(2)
    n, e, heads = g.number_of_nodes(), g.number_of_edges(), 4
    s = _leaf(e, heads, seed=4)
    s2 = s.detach().clone().requires_grad_(True)
    # reference formulation: exp / per-destination sum (stable form)
    mx = torch.full((n, heads), -float("inf"), device=device()).scatter_reduce(
        0, dst[:, None].expand(e, heads), s2.detach(), "amax")
    ex = torch.exp(s2 - mx[dst])
    want = ex / torch.zeros(n, heads, device=device()).index_add(0, dst, ex)[dst]
    got = dgl_b200.edge_softmax(g, s)
    close(got, want)
    gout = torch.randn(e, heads, generator=torch.Generator().manual_seed(5)).to(device())
    got.backward(gout)
    want.backward(gout)
    close(s.grad, s2.grad)
    # (E, H, 1) scores keep their shape
    assert dgl_b200.edge_softmax(g, s.detach().unsqueeze(-1)).shape == (e, heads, 1)


def test_gat_attention_matches_udf_formulation_with_gradients():
    """The GAT layer's edge path (softmax-normalised attention) three ways: composed kernels with
    autograd, the fused single launch (no grad), and per-edge gathers + torch autograd."""
    import dgl_b200
    g, src, dst = _setup(3, zero_deg_frac=0.0)
    n = g.number_of_nodes()
    # self loops so that no destination is empty (the example adds them, train.py:186-189)
    loops = torch.arange(n, device=device())
    g.add_edges(loops, loops)
    src, dst = torch.cat([src, loops]), torch.cat([dst, loops])
    e, heads, dim = g.number_of_edges(), 4, 6
    a1, a2, ft = _leaf(n, heads, 1, seed=6), _leaf(n, heads, 1, seed=7), _leaf(n, heads, dim, seed=8)
    b1, b2, ft2 = (t.detach().clone().requires_grad_(True) for t in (a1, a2, ft))
    s = torch.nn.functional.leaky_relu(b1[src] + b2[dst], 0.2).reshape(e, heads)
    mx = torch.full((n, heads), -float("inf"), device=device()).scatter_reduce(
        0, dst[:, None].expand(e, heads), s.detach(), "amax")
    ex = torch.exp(s - mx[dst])
    alpha_w = ex / torch.zeros(n, heads, device=device()).index_add(0, dst, ex)[dst]
    want = torch.zeros(n, heads, dim, device=device()).index_add(0, dst, alpha_w[:, :, None] * ft2[src])
    out, alpha = dgl_b200.gat_attention(g, a1, a2, ft, 0.2)
    close(out, want)
    close(alpha, alpha_w)
    gout = torch.randn(n, heads, dim, generator=torch.Generator().manual_seed(9)).to(device())
    out.backward(gout)
    want.backward(gout)
    close(ft.grad, ft2.grad)
    close(a1.grad, b1.grad)
    close(a2.grad, b2.grad)
    with torch.no_grad():
        fused, none = dgl_b200.gat_attention(g, a1, a2, ft, 0.2)
    assert none is None
    close(fused, want)


def test_gat_attention_slot_order_equals_edge_order():
    """The chain over Adjacency.slot_view() (scores / alpha in CSR slot order between the kernels) against the chain
    through the edge-id indirection: same output, same alpha in edge-id order, same gradients; multigraph with hubs."""
    import dgl_b200
    from tests.util import random_graph
    n = 300
    s_np, d_np = random_graph(n, 5000, 21, hubs=2, hub_deg=700, zero_deg_frac=0.1)
    g = dgl_b200.DGLGraph()
    g.add_nodes(n)
    g.add_edges(torch.from_numpy(s_np), torch.from_numpy(d_np))
    g.to(device())
    heads, dim = 4, 8
    res = []
    for slot in (False, True):
        a1, a2, ft = _leaf(n, heads, 1, seed=31), _leaf(n, heads, 1, seed=32), _leaf(n, heads, dim, seed=33)
        out, alpha = dgl_b200.gat_attention(g, a1, a2, ft, 0.2, slot_order=slot)
        gout = torch.randn(n, heads, dim, generator=torch.Generator().manual_seed(34)).to(device())
        (out * gout).sum().backward()
        res.append((out.detach(), alpha.detach(), a1.grad, a2.grad, ft.grad))
        out2, none = dgl_b200.gat_attention(g, a1, a2, ft, 0.2, want_alpha=False, slot_order=slot)
        assert none is None
        close(out2.detach(), out.detach())
    view, slot_of_edge = g.adjacency().slot_view()
    assert view.csr_in.eids is None and torch.equal(g.adjacency().csr_in.eids.long()[slot_of_edge],
                                                    torch.arange(len(s_np), device=device()))
    for a, b in zip(*res):
        close(a, b)
    # an adjacency with explicit edge ids (an edge subset) has no slot view: the chain falls back
    from dgl_b200.autograd import Adjacency
    sub = Adjacency(n, n, coo=(torch.from_numpy(d_np[:50]).to(device()), torch.from_numpy(s_np[:50]).to(device()),
                               torch.arange(50, device=device())))
    assert sub.slot_view() is None


def test_apply_edges_builtin_is_one_sddmm_executor():
    from dgl_b200 import function as fn
    g, src, dst = _setup(4)
    n = g.number_of_nodes()
    g.ndata["a1"], g.ndata["a2"] = _leaf(n, 4, 1, seed=10).detach(), _leaf(n, 4, 1, seed=11).detach()
    g.apply_edges(fn.u_add_v("a1", "a2", "a"))
    names = [ln.split("(")[0].split()[-1] for ln in g.last_prog.lines()]
    assert names == ["READ_COL", "READ_COL", "SDDMM", "WRITE_COL_"]
    want = g.ndata["a1"][src] + g.ndata["a2"][dst]
    close(g.edata["a"], want)
    # the same builtin called as a UDF (edge subset -> generic path) computes the same rows
    eids = torch.arange(0, g.number_of_edges(), 3, device=device())
    g.edata["b"] = torch.zeros_like(g.edata["a"])
    g.apply_edges(fn.u_add_v("a1", "a2", "b"), eids)
    close(g.edata["b"][eids], want[eids])
    assert float(g.edata["b"][1].abs().sum()) == 0.0
    g.apply_edges(fn.u_dot_v("a1", "a2", "d"))
    close(g.edata["d"], (g.ndata["a1"][src] * g.ndata["a2"][dst]).sum(-1, keepdim=True))
    g.apply_edges(fn.copy_u("a1", "c"))
    close(g.edata["c"], g.ndata["a1"][src])
