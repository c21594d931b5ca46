"""Generate tests/golden/gcn_layer_*.npz by running the UNMODIFIED reference's GCN example layer
(examples/pytorch/gcn/gcn_spmv.py:23-62, class GCNLayer) on its own DGLGraph.  TEST INFRASTRUCTURE ONLY;
build container only:

    DGL_LIBRARY_PATH=oracle/_ref PYTHONPATH=/root/reference/python python oracle/gen_golden_gcn.py

Stores inputs (edge list, features, weight, bias, norm), the layer's output with and without ReLU / bias,
and the gradients of sum(out * probe) with respect to the features, the weight and the bias.
"""
import importlib.util
import os
import sys

import numpy as np
import torch

import dgl

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")
EXAMPLE = "/root/reference/examples/pytorch/gcn/gcn_spmv.py"


def load_example():
    spec = importlib.util.spec_from_file_location("ref_gcn_spmv", EXAMPLE)
    mod = importlib.util.module_from_spec(spec)
    sys.argv = ["gcn_spmv.py"]
    spec.loader.exec_module(mod)
    return mod


def main():
    ex = load_example()
    for seed, (n, e, fin, fout) in enumerate([(60, 500, 12, 8), (150, 2500, 33, 16)]):
        rng = np.random.RandomState(100 + seed)
        src = rng.randint(0, n, e).astype(np.int64)
        dst = rng.randint(0, n - 6, e).astype(np.int64)
        dst[rng.choice(e, e // 6, replace=False)] = 2                 # hub; last 6 nodes have no in-edge
        src = np.concatenate([src, np.arange(n)])                      # self loops, gcn_spmv.py:135-136
        dst = np.concatenate([dst, np.arange(n)])
        g = dgl.DGLGraph()
        g.add_nodes(n)
        g.add_edges(torch.from_numpy(src), torch.from_numpy(dst))
        degs = g.in_degrees().float()
        norm = torch.pow(degs, -0.5)
        norm[torch.isinf(norm)] = 0
        g.ndata["norm"] = norm.unsqueeze(1)                            # gcn_spmv.py:138-143
        x = torch.from_numpy(rng.randn(n, fin).astype(np.float32))
        probe = torch.from_numpy(rng.randn(n, fout).astype(np.float32))
        rec = dict(n=n, src=src, dst=dst, x=x.numpy(), norm=norm.numpy(), probe=probe.numpy())
        for tag, act, bias in (("relu_bias", torch.relu, True), ("linear_nobias", None, False)):
            torch.manual_seed(7 + seed)
            layer = ex.GCNLayer(g, fin, fout, act, 0.0, bias=bias)
            xin = x.clone().requires_grad_(True)
            out = layer(xin)
            (out * probe).sum().backward()
            rec[tag + "_w"] = layer.weight.detach().numpy()
            rec[tag + "_out"] = out.detach().numpy()
            rec[tag + "_dx"] = xin.grad.numpy()
            rec[tag + "_dw"] = layer.weight.grad.numpy()
            if bias:
                rec[tag + "_b"] = layer.bias.detach().numpy()
                rec[tag + "_db"] = layer.bias.grad.numpy()
        np.savez_compressed(os.path.join(OUT, "gcn_layer_%d.npz" % seed), **rec)
        print("wrote gcn_layer_%d.npz" % seed)


if __name__ == "__main__":
    main()
