"""TEST INFRASTRUCTURE ONLY - a CPU stand-in for `dgl_b200.ops` so that the HOST layer (graph.py,
scheduler.py, autograd.py, dist.py: id plumbing, plans, write-back, message frames) can be
exercised by `-m "not gpu"` tests in a container without a GPU.

It restates the documented semantics of every C-ABI entry point the host layer calls
(include/dgl_b200.h) with plain torch CPU ops, in the spirit of oracle/: slow, obvious, small cases
only.  The product never imports this file; `dgl_b200.ops` itself has no CPU path and keeps raising
DGLError on CPU tensors outside of the `installed()` context.  What these tests prove is the host
logic, not the kernels - kernel parity is the `-m gpu` suite.
"""
import contextlib

import torch

from dgl_b200 import ops
from dgl_b200 import graph as graph_mod


def _csr_build(row, col, n_rows, n_cols, eid=None, order=0):
    nnz = int(row.shape[0])
    assert row.dtype == torch.int64 and col.dtype == torch.int64
    if nnz and (int(row.min()) < 0 or int(row.max()) >= n_rows or int(col.min()) < 0 or int(col.max()) >= n_cols):
        raise ops.DGLError("csr_build: id out of range")
    pos = torch.arange(nnz, dtype=torch.int64)
    if order == 0:
        perm = torch.sort(row, stable=True)[1]
    else:
        perm = torch.sort(row * max(n_cols, 1) + col, stable=True)[1]
    indptr = torch.zeros(n_rows + 1, dtype=torch.int64)
    indptr[1:] = torch.cumsum(torch.bincount(row, minlength=n_rows), 0)
    e = pos if eid is None else eid
    return ops.CSR(n_rows, n_cols, indptr, col[perm].to(torch.int32), e[perm].to(torch.int32))


def _widen(t):
    return t.to(torch.int64)


def _degrees(csr):
    return csr.indptr[1:] - csr.indptr[:-1]


def _csr_row_ids(csr):
    return torch.repeat_interleave(torch.arange(csr.n_rows, dtype=torch.int64), _degrees(csr))


def _csr_select(csr, rows, want=("which", "slot")):
    if rows.numel() and (int(rows.min()) < 0 or int(rows.max()) >= csr.n_rows):
        raise ops.DGLError("Invalid vertex")
    lo = csr.indptr[rows]
    cnt = csr.indptr[rows + 1] - lo
    total = int(cnt.sum())
    which = torch.repeat_interleave(torch.arange(rows.numel()), cnt, output_size=total)
    start = torch.cumsum(cnt, 0) - cnt
    slot = lo[which] + (torch.arange(total) - start[which])
    outs = {"which": which, "slot": slot, "row": rows[which], "col": csr.indices[slot].long(),
            "eid": csr.eids[slot].long() if csr.eids is not None else slot}
    return tuple(outs[k] for k in want)


def _degrees_of(csr, rows):
    return csr.indptr[rows + 1] - csr.indptr[rows]


def _slot_edge(csr):
    return csr.eids.long() if csr.eids is not None else torch.arange(csr.nnz, dtype=torch.int64)


def _messages(csr, msg_op, x, w, w_width):
    """(nnz, feat) fp32 messages in slot order."""
    e = _slot_edge(csr)
    if msg_op == ops.MSG_COPY_SRC:
        return x[csr.indices.long()].float()
    if msg_op == ops.MSG_COPY_EDGE:
        return w[e].float()
    feat = x.shape[1]
    wv = w.reshape(w.shape[0], -1)[e].float()
    return x[csr.indices.long()].float() * wv.repeat_interleave(feat // w_width, dim=1)


def _spmm(csr, msg_op, reduce_op, x=None, w=None, w_width=1, want_arg=False, chunk=0, out=None):
    ref = x if x is not None else w
    if msg_op == ops.MSG_COPY_EDGE:
        w_width = int(w.shape[1])
    m = _messages(csr, msg_op, x, w, w_width)
    feat = m.shape[1]
    rows = _csr_row_ids(csr)
    e = _slot_edge(csr)
    arg = None
    if reduce_op == ops.RED_SUM:
        y = torch.zeros((csr.n_rows, feat), dtype=torch.float32)
        for p in range(csr.nnz):          # slot order == the reference's COO accumulation order
            y[rows[p]] += m[p]
    else:
        y = torch.zeros((csr.n_rows, feat), dtype=torch.float32)
        a = torch.full((csr.n_rows, feat), -1, dtype=torch.int64)
        for p in range(csr.nnz):
            r = rows[p]
            better = (a[r] < 0) | (m[p] > y[r])
            y[r] = torch.where(better, m[p], y[r])
            a[r] = torch.where(better, e[p], a[r])
        if want_arg:
            arg = a.to(torch.int32)
    y = y.to(ref.dtype)
    if out is not None:
        out.copy_(y)
        y = out
    return y, arg


def _spmm_backward_x(csr_t, dy, w=None, w_width=1, arg=None, chunk=0):
    feat = dy.shape[1]
    e = _slot_edge(csr_t)
    g = dy[csr_t.indices.long()].float()
    if w is not None:
        g = g * w.reshape(w.shape[0], -1)[e].float().repeat_interleave(feat // w_width, dim=1)
    if arg is not None:
        g = g * (arg[csr_t.indices.long()].long() == e[:, None]).float()
    dx = torch.zeros((csr_t.n_rows, feat), dtype=torch.float32)
    dx.index_add_(0, _csr_row_ids(csr_t), g)
    return dx.to(dy.dtype)


def _sddmm(csr, op, u, v, heads, dim):
    rows = _csr_row_ids(csr)
    s = csr.indices.long()
    e = _slot_edge(csr)
    if op == ops.SDDMM_DOT:
        val = (u[s].reshape(-1, heads, dim).float() * v[rows].reshape(-1, heads, dim).float()).sum(-1)
    elif op == ops.SDDMM_ADD:
        val = u[s].float() + v[rows].float()
    elif op == ops.SDDMM_MUL:
        val = u[s].float() * v[rows].float()
    elif op == ops.SDDMM_COPY_U:
        val = u[s].float()
    else:
        val = v[rows].float()
    ref = u if u is not None else v
    out = torch.empty((csr.nnz, val.shape[1]), dtype=ref.dtype)
    out[e] = val.reshape(csr.nnz, -1).to(ref.dtype)
    return out


def _unique_sorted(ids, id_bound, want_inverse=True):
    uniq, inv = torch.unique(ids, sorted=True, return_inverse=True)
    return uniq, (inv if want_inverse else None)


def _degree_bucketing(msg_ids, vids, recv_ids, id_bound):
    from oracle import oracle
    res = oracle.degree_bucketing(msg_ids.numpy(), vids.numpy(), recv_ids.numpy())
    return tuple(torch.from_numpy(a) for a in res)


def _gather_rows(x, idx, out=None, max_ctas=0):
    val = x if idx is None else x[idx]
    if out is None:
        return val.clone()
    out.copy_(val.reshape(out.shape))
    return out


def _edge_data_to_slots(csr, w):
    return w if csr.eids is None else w[csr.eids.long()]


def _edge_softmax(csr, score):
    rows = _csr_row_ids(csr)
    e = _slot_edge(csr)
    s = score.float()[e]
    alpha = torch.zeros_like(score, dtype=torch.float32)
    for r in range(csr.n_rows):
        sel = (rows == r).nonzero().reshape(-1)
        if sel.numel():
            alpha[e[sel]] = torch.softmax(s[sel], 0)
    return alpha


def _edge_softmax_backward(csr, alpha, dalpha):
    rows = torch.empty(csr.nnz, dtype=torch.int64)
    rows[_slot_edge(csr)] = _csr_row_ids(csr)          # destination of every edge id
    dot = torch.zeros((csr.n_rows, alpha.shape[1]), dtype=torch.float32)
    dot.index_add_(0, rows, alpha * dalpha)
    return alpha * (dalpha - dot[rows])


def _gat_aggregate(csr, a_src, a_dst, ft, negative_slope, flavour=1, want_edge=True):
    heads, dim = int(ft.shape[1]), int(ft.shape[2])
    rows = _csr_row_ids(csr)
    s_idx = csr.indices.long()
    e = _slot_edge(csr)
    s = torch.nn.functional.leaky_relu(a_src.reshape(-1, heads)[s_idx] + a_dst.reshape(-1, heads)[rows], negative_slope)
    if flavour == ops.SOFTMAX_GAT_EXAMPLE:
        a = torch.exp(s).clamp(-10, 10)
    else:
        mx = torch.full((csr.n_rows, heads), -float("inf")).scatter_reduce(0, rows[:, None].expand_as(s), s, "amax")
        a = torch.exp(s - mx[rows])
    z = torch.zeros((csr.n_rows, heads)).index_add_(0, rows, a)
    num = torch.zeros((csr.n_rows, heads, dim)).index_add_(0, rows, a[:, :, None] * ft[s_idx])
    out = torch.where(z[:, :, None] > 0, num / z[:, :, None].clamp_min(1e-38), torch.zeros_like(num))
    a_out = None
    if want_edge:
        a_out = torch.zeros((csr.nnz, heads))
        a_out[e] = a
    return out, a_out, z


_PATCHES = {
    "csr_build": _csr_build, "widen": _widen, "degrees": _degrees, "csr_row_ids": _csr_row_ids,
    "spmm": _spmm, "spmm_backward_x": _spmm_backward_x, "sddmm": _sddmm,
    "unique_sorted": _unique_sorted, "degree_bucketing": _degree_bucketing,
    "gather_rows": _gather_rows, "edge_data_to_slots": _edge_data_to_slots,
    "edge_softmax": _edge_softmax, "edge_softmax_backward": _edge_softmax_backward,
    "gat_aggregate": _gat_aggregate, "csr_select": _csr_select, "degrees_of": _degrees_of,
}


@contextlib.contextmanager
def installed():
    """Swap the kernel-launching functions of dgl_b200.ops for the CPU restatements above and let
    DGLGraph accept a CPU device, for the duration of a test."""
    saved = {k: getattr(ops, k) for k in _PATCHES}
    saved_need = graph_mod.DGLGraph._need_device
    if not hasattr(ops, "DGLError"):
        from dgl_b200.base import DGLError
        ops.DGLError = DGLError
    try:
        for k, fn in _PATCHES.items():
            setattr(ops, k, fn)
        graph_mod.DGLGraph._need_device = lambda self: self._device or torch.device("cpu")
        yield
    finally:
        for k, fn in saved.items():
            setattr(ops, k, fn)
        graph_mod.DGLGraph._need_device = saved_need
