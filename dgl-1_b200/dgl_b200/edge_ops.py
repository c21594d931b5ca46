"""Differentiable per-edge operators over an Adjacency: SDDMM (u-op-v), per-destination edge
softmax, and the GAT attention aggregation composed from them.

This is the edge side of the path.  The reference reaches it through
apply_edges + a Python edge UDF and gets gradients from torch autograd over three E-row gathers
(examples/pytorch/gat/train.py:61-97; runtime/scheduler.py:234-277, 572-583).  Here every forward AND
every gradient is one kernel of libdgl_b200.so:

  op                 forward                           d/du                           d/dv
  u_add_v            sddmm ADD                         copy_edge->sum over out-CSR    copy_edge->sum over in-CSR
  u_mul_v            sddmm MUL                         (v * g)->sum over out-CSR      (u * g)->sum over in-CSR
  u_dot_v            sddmm DOT                         (v * g_head)->sum, out-CSR     (u * g_head)->sum, in-CSR
  copy_u / copy_v    sddmm COPY_U / COPY_V             copy_edge->sum                 copy_edge->sum
  edge_softmax       edge_softmax_csr                  edge_softmax_backward_csr

"(v * g)->sum over the out-CSR" is kernel (1) in its SRC_MUL_EDGE form on the CSR keyed by source,
whose "source" operand is then v (indexed by the slot's destination) - the same trick the gradient of
update_all uses.  No E x D intermediate is materialised beyond the op's own output.
"""
import torch

from . import ops

_OPS = {"u_dot_v": ops.SDDMM_DOT, "u_add_v": ops.SDDMM_ADD, "u_mul_v": ops.SDDMM_MUL,
        "copy_u": ops.SDDMM_COPY_U, "copy_v": ops.SDDMM_COPY_V}


def _adj(g):
    """Accept a DGLGraph (its cached whole-graph adjacency) or an Adjacency."""
    return g.adjacency() if hasattr(g, "adjacency") else g


def _sum_edges(csr, g):
    """Per-row sum of edge rows g (E, F) indexed by edge id: copy_edge -> sum."""
    return ops.spmm(csr, ops.MSG_COPY_EDGE, ops.RED_SUM, w=g)[0]


def _sum_weighted(csr, x, g, w_width):
    """Per-row sum of x[other endpoint] * g[edge]: src_mul_edge -> sum with edge rows g."""
    return ops.spmm(csr, ops.MSG_SRC_MUL_EDGE, ops.RED_SUM, x=x, w=g, w_width=w_width)[0]


class _SDDMM(torch.autograd.Function):
    @staticmethod
    def forward(ctx, adj, op, u, v, heads, dim):
        ctx.adj, ctx.op, ctx.heads, ctx.dim = adj, op, heads, dim
        ctx.save_for_backward(u, v)
        return ops.sddmm(adj.csr_in, op, u, v, heads, dim)

    @staticmethod
    def backward(ctx, g):
        u, v = ctx.saved_tensors
        adj, op, heads, dim = ctx.adj, ctx.op, ctx.heads, ctx.dim
        g = g.contiguous()
        du = dv = None
        want_u = u is not None and ctx.needs_input_grad[2]
        want_v = v is not None and ctx.needs_input_grad[3]
        if op in (ops.SDDMM_ADD, ops.SDDMM_COPY_U, ops.SDDMM_COPY_V):
            if want_u and op != ops.SDDMM_COPY_V:
                du = _sum_edges(adj.csr_out, g)
            if want_v and op != ops.SDDMM_COPY_U:
                dv = _sum_edges(adj.csr_in, g)
        else:
            # MUL: one weight per element; DOT: one weight per head
            w_width = heads * dim if op == ops.SDDMM_MUL else heads
            if want_u:
                du = _sum_weighted(adj.csr_out, v, g, w_width)
            if want_v:
                dv = _sum_weighted(adj.csr_in, u, g, w_width)
        return None, None, du, dv, None, None


def sddmm(adj, op, u=None, v=None):
    """Per-edge u-op-v (op: 'u_add_v' | 'u_mul_v' | 'u_dot_v' | 'copy_u' | 'copy_v').

    u: (n_src, H, D) or (n_src, F); v: (n_dst, ...) of the same trailing shape.  Returns (E, ...)
    indexed by edge id: the trailing shape of the operands, or (..., 1) for u_dot_v (the dot runs
    over the last dim).  Differentiable in u and v."""
    adj = _adj(adj)
    code = _OPS[op]
    ref = u if u is not None else v
    tail = tuple(ref.shape[1:])
    if code == ops.SDDMM_DOT:
        dim = int(tail[-1]) if tail else 1
        heads = 1
        for t in tail[:-1]:
            heads *= int(t)
        out_tail = tail[:-1] + (1,)
    else:
        heads, dim = 1, 1
        for t in tail:
            dim *= int(t)
        out_tail = tail
    flat = lambda t: None if t is None else t.reshape(t.shape[0], -1)  # noqa: E731
    out = _SDDMM.apply(adj, code, flat(u), flat(v), heads, dim)
    return out.reshape((out.shape[0],) + out_tail)


class _EdgeSoftmax(torch.autograd.Function):
    @staticmethod
    def forward(ctx, adj, score):
        alpha = ops.edge_softmax(adj.csr_in, score)
        ctx.adj = adj
        ctx.save_for_backward(alpha)
        return alpha

    @staticmethod
    def backward(ctx, g):
        (alpha,) = ctx.saved_tensors
        return None, ops.edge_softmax_backward(ctx.adj.csr_in, alpha, g.contiguous())


def edge_softmax(adj, score):
    """Softmax of edge scores (E, H[, 1]) over the in-edges of every destination, per head
    (numerically stable; fp32).  Differentiable."""
    shape = score.shape
    adj = _adj(adj)
    return _EdgeSoftmax.apply(adj, score.reshape(shape[0], -1).float()).reshape(shape)


def gat_attention(adj, a_src, a_dst, ft, negative_slope=0.2, want_alpha=None, slot_order=True):
    """GAT edge path with a softmax-normalised attention, end to end on kernels and differentiable:
    alpha = softmax_dst(leaky_relu(a_src[src] + a_dst[dst])),  out[v] = sum_e alpha_e * ft[src_e].
    a_src: (n_src, H[, 1]), a_dst: (n_dst, H[, 1]), ft: (n_src, H, D).  Returns (out (n_dst, H, D),
    alpha (E, H) in edge-id order, or None).  When no gradient is needed and alpha is not asked for, the whole
    thing is ONE launch (dglb200_gat_aggregate_csr, stable flavour); otherwise SDDMM -> leaky_relu ->
    edge softmax -> weighted SpMM, each with its gradient kernel.  want_alpha=False skips the copy of alpha
    back to edge-id order on that path.
    slot_order: run the chain over Adjacency.slot_view() - scores and alpha live in CSR slot order between the
    kernels, so none of them goes through the edge-id indirection."""
    from .autograd import aggregate
    adj = _adj(adj)
    heads, dim = int(ft.shape[1]), int(ft.shape[2])
    a_src = a_src.reshape(a_src.shape[0], heads)
    a_dst = a_dst.reshape(a_dst.shape[0], heads)
    need_grad = torch.is_grad_enabled() and (a_src.requires_grad or a_dst.requires_grad or ft.requires_grad)
    if not need_grad and not want_alpha:
        out, _, _ = ops.gat_aggregate(adj.csr_in, a_src, a_dst, ft, negative_slope,
                                      flavour=ops.SOFTMAX_STABLE, want_edge=False)
        return out, None
    view = adj.slot_view() if slot_order else None
    work, slot_of_edge = view if view is not None else (adj, None)
    score = torch.nn.functional.leaky_relu(sddmm(work, "u_add_v", a_src, a_dst), negative_slope)
    alpha = edge_softmax(work, score)
    out = aggregate(work, ops.MSG_SRC_MUL_EDGE, ops.RED_SUM, x=ft.reshape(ft.shape[0], heads * dim),
                    w=alpha, w_width=heads)
    if want_alpha is False:
        alpha = None
    elif slot_of_edge is not None:
        alpha = alpha[slot_of_edge]
    return out.reshape(-1, heads, dim), alpha
