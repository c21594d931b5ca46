"""-m gpu: the fused GCN layer (dgl_b200.fused) against golden vectors produced by the UNMODIFIED reference's
example layer (examples/pytorch/gcn/gcn_spmv.py:23-62 on the reference's own DGLGraph; generator:
oracle/gen_golden_gcn.py): output with ReLU + bias and without, and the gradients of the features, the
weight and the bias.  Tolerance: the reference's own rtol = atol = 1e-4 (tests/backend/pytorch/__init__.py:11-12)
tightened to 1e-5 relative to the largest value."""
import glob
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "gcn_layer_*.npz")))


def device():
    return torch.device("cuda")


def _close(got, want, rtol=1e-5):
    got = got.detach().cpu().numpy()
    np.testing.assert_allclose(got, want, rtol=rtol, atol=rtol * (np.abs(want).max() + 1e-30))


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_fused_layer_matches_reference_example_layer(path):
    import dgl_b200
    from dgl_b200 import fused
    c = dict(np.load(path))
    dev = device()
    g = dgl_b200.DGLGraph()
    g.add_nodes(int(c["n"]))
    g.add_edges(torch.from_numpy(c["src"]), torch.from_numpy(c["dst"]))
    g.to(dev)
    norm = torch.from_numpy(c["norm"]).to(dev)
    probe = torch.from_numpy(c["probe"]).to(dev)
    for tag, act, has_bias in (("relu_bias", "relu", True), ("linear_nobias", None, False)):
        x = torch.from_numpy(c["x"]).to(dev).requires_grad_(True)
        w = torch.from_numpy(c[tag + "_w"]).to(dev).requires_grad_(True)
        b = torch.from_numpy(c[tag + "_b"]).to(dev).requires_grad_(True) if has_bias else None
        out = fused.gcn_layer(g, x, w, bias=b, norm=norm, activation=act)
        _close(out, c[tag + "_out"])
        (out * probe).sum().backward()
        _close(x.grad, c[tag + "_dx"])
        _close(w.grad, c[tag + "_dw"])
        if has_bias:
            _close(b.grad, c[tag + "_db"])
        # the unfused composition through the public API gives the same numbers
        from dgl_b200 import function as fn
        g.ndata["h"] = torch.mm(x.detach(), w.detach()) * norm.unsqueeze(1)
        g.update_all(fn.copy_src("h", "m"), fn.sum("m", "h"))
        ref = g.ndata.pop("h") * norm.unsqueeze(1)
        if has_bias:
            ref = torch.relu(ref + b.detach())
        _close(out, ref.cpu().numpy())


def test_epilogue_on_long_rows_ragged_width_and_empty_rows():
    """row_scale / bias / relu in the store epilogue AND in the long-row fix-up kernel, on a ragged width."""
    from dgl_b200 import ops
    from oracle import oracle
    from tests.util import random_graph
    n, feat = 300, 41
    src, dst = random_graph(n, 6000, 3, hubs=2, hub_deg=900, zero_deg_frac=0.2)
    rng = np.random.RandomState(0)
    x = rng.randn(n, feat).astype(np.float32)
    rs = rng.rand(n).astype(np.float32)
    bias = rng.randn(feat).astype(np.float32)
    csr = ops.csr_build(torch.from_numpy(dst).to(device()), torch.from_numpy(src).to(device()), n, n)
    want = np.maximum(oracle.update_all(src, dst, n, "copy_src", "sum", x=x) * rs[:, None] + bias, 0)
    for chunk in (0, 64):
        y, _ = ops.spmm(csr, ops.MSG_COPY_SRC, ops.RED_SUM, x=torch.from_numpy(x).to(device()), chunk=chunk,
                        row_scale=torch.from_numpy(rs).to(device()), bias=torch.from_numpy(bias).to(device()), relu=True)
        _close(y, want)
    with pytest.raises(Exception):
        ops.spmm(csr, ops.MSG_COPY_SRC, ops.RED_MAX, x=torch.from_numpy(x).to(device()), relu=True)
