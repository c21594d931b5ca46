"""-m gpu: a whole training step of the builtin path - DGLGraph.update_all forward, its gradient kernels and an SGD
update - captured into ONE CUDA graph and replayed (the Cora / PubMed / AIFB-sized configs are
launch-latency-bound).  What this pins: once the structure (CSR pair, long-row plan) exists, a step makes no
size-returning call, no stream synchronisation and no cudaMalloc - any of them would abort the capture.
Collected late (a failed capture leaves the process's CUDA state unusable); numbers: profiles/r02_graph_capture.jsonl.
"""
import numpy as np
import pytest
import torch

from tests.util import random_graph

pytestmark = pytest.mark.gpu


def _capture(step, warm=3):
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(warm):
            step()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        out = step()
    return g, out


@pytest.mark.parametrize("kind", ["copy_src_sum", "src_mul_edge_sum", "hub_rows"])
def test_training_step_replays_from_one_cuda_graph(kind):
    import dgl_b200
    import dgl_b200.function as fn
    from dgl_b200 import _lib
    dev = torch.device("cuda")
    n, feat = 700, 16
    hubs = 3 if kind == "hub_rows" else 0                   # hub rows: the long-row slices + fix-up kernel too
    src, dst = random_graph(n, 9000, 5, hubs=hubs, hub_deg=1500, zero_deg_frac=0.1)
    g = dgl_b200.DGLGraph()
    g.add_nodes(n)
    g.add_edges(torch.from_numpy(src), torch.from_numpy(dst))
    g.to(dev)
    gen = torch.Generator(device="cpu").manual_seed(3)
    x = torch.randn(n, 32, generator=gen).to(dev)
    w = (torch.randn(32, feat, generator=gen) * 0.1).to(dev).requires_grad_(True)
    b = torch.zeros(feat, device=dev, requires_grad=True)
    ew = torch.rand(len(src), 1, generator=gen).to(dev).requires_grad_(True)
    params = [w, b] + ([ew] if kind == "src_mul_edge_sum" else [])

    def step():
        g.ndata["h"] = torch.mm(x, w)
        if kind == "src_mul_edge_sum":
            g.edata["w"] = ew
            g.update_all(fn.src_mul_edge("h", "w", "m"), fn.sum("m", "h"))
        else:
            g.update_all(fn.copy_src("h", "m"), fn.sum("m", "h"))
        loss = (torch.relu(g.ndata.pop("h") + b)).square().mean()
        loss.backward()
        with torch.no_grad():
            for p in params:
                p.add_(p.grad, alpha=-0.05)
                p.grad = None
        return loss

    graph, loss = _capture(step)                            # capture FIRST (AccumulateGrad nodes on the side stream)
    start = [p.detach().clone() for p in params]
    l0 = _lib.launch_count()
    graph.replay()
    torch.cuda.synchronize()
    assert _lib.launch_count() == l0                        # a replay goes through no entry point of the library
    replayed = [p.detach().clone() for p in params]
    replay_loss = float(loss)
    with torch.no_grad():
        for p, q in zip(params, start):
            p.copy_(q)
    eager_loss = float(step())
    torch.cuda.synchronize()
    assert _lib.launch_count() > l0
    assert np.isfinite(replay_loss) and abs(replay_loss - eager_loss) <= 1e-5 * abs(eager_loss) + 1e-7
    assert not torch.equal(replayed[0], start[0])            # the replay did update the weights
    for a, p in zip(replayed, params):
        assert torch.allclose(a, p.detach(), rtol=1e-5, atol=1e-6)
    # several replays keep stepping the weights (the graph reads and writes the live parameter tensors)
    before = w.detach().clone()
    for _ in range(3):
        graph.replay()
    torch.cuda.synchronize()
    assert not torch.equal(before, w.detach())
