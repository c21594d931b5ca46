"""Fused GCN layer: the reference example's layer
(examples/pytorch/gcn/gcn_spmv.py:45-62)

    h = h @ W;  h = h * norm;  update_all(copy_src, sum);  h = h * norm;  h = h + bias;  h = act(h)

runs four eager element-wise passes over N x F around every aggregation.  Here the destination-side
normalisation, the bias and the ReLU live in the STORE EPILOGUE of the SpMM kernel
(dglb200_spmm_csr_ext: row_scale / bias / relu), and in the backward the source-side normalisation is
the epilogue of the transposed SpMM.  What stays a separate pass: the source-side scaling of the GEMM
output in the forward (folding it into the gather would add a 4-byte gather + FFMA per edge to a loop
that is L2-bound at mean degree 492: the scalar-weight kernel measures 8.2 ms against 6.2 ms per SpMM at
C2, while the pass over N x F costs 0.1 ms) and the ReLU mask x destination norm of the incoming gradient.
"""
import torch

from . import ops


class _GCNAggregate(torch.autograd.Function):
    """y = act(dnorm * (A (snorm * hw)) + bias); gradients for hw and bias."""

    @staticmethod
    def forward(ctx, adj, hw, snorm, dnorm, bias, relu):
        hs = hw * snorm.unsqueeze(1)                       # source-side normalisation (one pass)
        y, _ = ops.spmm(adj.csr_in, ops.MSG_COPY_SRC, ops.RED_SUM, x=hs.contiguous(),
                        row_scale=dnorm, bias=bias, relu=relu)
        ctx.adj, ctx.relu, ctx.has_bias = adj, relu, bias is not None
        ctx.save_for_backward(y if relu else None, snorm, dnorm)
        return y

    @staticmethod
    def backward(ctx, dy):
        y, snorm, dnorm = ctx.saved_tensors
        g = dy * (y > 0) if ctx.relu else dy
        dbias = g.sum(0) if ctx.has_bias and ctx.needs_input_grad[4] else None
        dhw = None
        if ctx.needs_input_grad[1]:
            g = (g * dnorm.unsqueeze(1)).contiguous()      # destination-side normalisation of the gradient
            # d(hw) = snorm * (A^T g): the source-side normalisation is the epilogue of the transposed SpMM
            dhw, _ = ops.spmm(ctx.adj.csr_out, ops.MSG_COPY_SRC, ops.RED_SUM, x=g, row_scale=snorm)
        return None, dhw, None, None, dbias, None


def gcn_aggregate(adj, hw, snorm, dnorm=None, bias=None, relu=False):
    """act(dnorm[v] * sum_{u -> v} snorm[u] * hw[u] + bias) for every destination v, differentiable in hw and
    bias.  adj: autograd.Adjacency (g.adjacency()); snorm / dnorm: (N,) float32 (dnorm defaults to snorm,
    the symmetric normalisation of the example)."""
    snorm = snorm.reshape(-1).contiguous().float()
    dnorm = snorm if dnorm is None else dnorm.reshape(-1).contiguous().float()
    return _GCNAggregate.apply(adj, hw, snorm, dnorm, bias, bool(relu))


def gcn_layer(g, h, weight, bias=None, norm=None, activation=None):
    """The example's GCNLayer.forward (dropout excluded) on a dgl_b200.DGLGraph: dense GEMM (torch / cuBLAS,
    as in the reference) + the fused aggregation.  activation: None or "relu" / torch.relu (fused); any
    other callable is applied afterwards."""
    if norm is None:
        deg = g.in_degrees().float()
        norm = deg.pow(-0.5)
        norm[torch.isinf(norm)] = 0
    fuse_relu = activation in ("relu", torch.relu, torch.nn.functional.relu)
    out = gcn_aggregate(g.adjacency(), torch.mm(h, weight), norm, bias=bias, relu=fuse_relu)
    if activation is not None and not fuse_relu:
        out = activation(out)
    return out
