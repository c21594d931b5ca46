"""Differentiable aggregation: the engine's replacement for torch's `sparse.mm` autograd.

In the reference the result of SPMVExecutor.run carries th.sparse.mm's backward
(python/dgl/backend/pytorch/tensor.py:143-146): d/dB = A^T * grad, and for
SPMV_WITH_DATA also d/dw_e = <B[src_e], grad[dst_e]>.  Here forward is kernel (1) on the CSR keyed
by destination, d/dx is kernel (2) (the same kernel on the CSR keyed by source), d/dw is the
SDDMM kernel (3).  max routes gradients through the argmax edge ids the forward kernel records.
"""
import torch

from . import ops


class Adjacency(object):
    """The pair of CSRs one aggregation needs: `rows` (keyed by destination, forward) and, built
    on first use, `cols` (keyed by source, gradient).  Plays the role of the cached sparse
    adjacency of GraphIndex.adjacency_matrix (python/dgl/graph_index.py:537-585) and of
    build_adj_matrix_uv for edge subsets (python/dgl/runtime/spmv.py:195-227)."""

    def __init__(self, n_dst, n_src, csr_in=None, csr_out=None, coo=None):
        self.n_dst = int(n_dst)
        self.n_src = int(n_src)
        self._in = csr_in
        self._out = csr_out
        self._coo = coo          # (dst_row, src_col, eid or None) int64 CUDA tensors
        # edge ids are 0..nnz-1 (a whole graph built from its edge list): the precondition of slot_view()
        self._dense_eids = coo is not None and coo[2] is None
        self._slot = None

    @property
    def csr_in(self):
        if self._in is None:
            r, c, e = self._coo
            self._in = ops.csr_build(r, c, self.n_dst, self.n_src, eid=e)
        return self._in

    @property
    def csr_out(self):
        if self._out is None:
            r, c, e = self._coo
            self._out = ops.csr_build(c, r, self.n_src, self.n_dst, eid=e)
        return self._out

    def slot_view(self):
        """(Adjacency, slot_of_edge) - the same graph with its edges RENUMBERED by their slot in csr_in, or None
        when the edge ids are not 0..nnz-1.  Edge data produced and consumed by a chain of kernels over this view
        (SDDMM -> edge softmax -> weighted SpMM and their gradients) is then read and written sequentially by the
        row-wise kernels instead of through the edge-id indirection (random 4 * H-byte accesses on a graph built
        from an unsorted edge list); slot_of_edge[e] maps a result back to the caller's edge order."""
        if not self._dense_eids:
            return None
        if self._slot is None:
            cin, cout = self.csr_in, self.csr_out
            dev = cin.device
            slot_of_edge = torch.empty(cin.nnz, dtype=torch.int64, device=dev)
            slot_of_edge[cin.eids.long()] = torch.arange(cin.nnz, dtype=torch.int64, device=dev)
            vin = ops.CSR(cin.n_rows, cin.n_cols, cin.indptr, cin.indices, None)
            vout = ops.CSR(cout.n_rows, cout.n_cols, cout.indptr, cout.indices, slot_of_edge[cout.eids.long()].int())
            view = Adjacency(self.n_dst, self.n_src, csr_in=vin, csr_out=vout)
            self._slot = (view, slot_of_edge)
        return self._slot

    def drop_coo(self):
        _ = self.csr_in, self.csr_out
        self._coo = None

    @property
    def nnz(self):
        return self.csr_in.nnz


class _Aggregate(torch.autograd.Function):
    @staticmethod
    def forward(ctx, adj, msg_op, reduce_op, w_width, x, w):
        need_grad = (x is not None and x.requires_grad) or (w is not None and w.requires_grad)
        want_arg = reduce_op == ops.RED_MAX and need_grad
        y, arg = ops.spmm(adj.csr_in, msg_op, reduce_op, x=x, w=w, w_width=w_width,
                          want_arg=want_arg)
        ctx.adj, ctx.msg_op, ctx.reduce_op, ctx.w_width = adj, msg_op, reduce_op, w_width
        ctx.save_for_backward(x, w, arg)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, arg = ctx.saved_tensors
        adj, msg_op, w_width = ctx.adj, ctx.msg_op, ctx.w_width
        dy = dy.contiguous()
        feat = dy.shape[1]
        dx = dw = None
        is_max = ctx.reduce_op == ops.RED_MAX
        if msg_op != ops.MSG_COPY_EDGE and ctx.needs_input_grad[4]:
            dx = ops.spmm_backward_x(adj.csr_out, dy,
                                     w=w if msg_op == ops.MSG_SRC_MUL_EDGE else None,
                                     w_width=w_width, arg=arg if is_max else None)
        if w is not None and ctx.needs_input_grad[5]:
            if not is_max:
                if msg_op == ops.MSG_COPY_EDGE:
                    dw = ops.sddmm(adj.csr_in, ops.SDDMM_COPY_V, None, dy, 1, feat)
                else:
                    dw = ops.sddmm(adj.csr_in, ops.SDDMM_DOT, x, dy, w_width, feat // w_width)
            else:
                dw = _max_edge_grad(adj, msg_op, x, w, w_width, arg, dy)
        return None, None, None, None, dx, dw


def _max_edge_grad(adj, msg_op, x, w, w_width, arg, dy):
    """d/dw under max: only the winning edge of every (row, feature) receives gradient.  Rare
    path (the reference reaches it through degree-bucketing autograd); index plumbing in torch."""
    feat = dy.shape[1]
    win = arg >= 0
    rows, cols = win.nonzero(as_tuple=True)
    e = arg[rows, cols].long()
    g = dy[rows, cols]
    if msg_op == ops.MSG_SRC_MUL_EDGE:
        # the source of edge e: slot -> (eid, src) through the in-edge CSR
        csr = adj.csr_in
        src_of_edge = torch.empty(csr.nnz, dtype=torch.long, device=dy.device)
        src_of_edge[csr.eids.long()] = csr.indices.long()
        g = g * x[src_of_edge[e], cols]
    dw = torch.zeros_like(w)
    grp = cols // (feat // w_width)
    dw.index_put_((e, grp), g, accumulate=True)
    return dw


def aggregate(adj, msg_op, reduce_op, x=None, w=None, w_width=1):
    """y (n_dst, feat) = reduce over in-edges of msg(x[src], w[eid]); differentiable in x and w."""
    return _Aggregate.apply(adj, msg_op, reduce_op, w_width, x, w)
