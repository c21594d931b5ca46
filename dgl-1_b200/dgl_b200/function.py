"""Builtin message / reduce functions: the drop-in surface of `dgl.function`.

Same factories, class names and public fields as the reference
(python/dgl/function/message.py:47,102,145,190-257; reducer.py:31-97; base.py:13-38):
`copy_src`, `src_mul_edge`, `copy_edge` -> `sum`, `max`.  Every builtin still carries its
UDF-equivalent `__call__` (the semantic spec, message.py:75-91 / reducer.py:45-46), which the
scheduler uses only when a (message, reduce) pair cannot be lowered to one kernel.

Difference from the reference: `is_spmv_supported` is wider.  The reference lowers only
sum with scalar edge weights (message.py:37-44, reducer.py:40-43); here max, copy_edge and
per-head / element-wise weights are kernels too (see `kernel_form`).
"""
import builtins
import operator

import torch

__all__ = ["src_mul_edge", "copy_src", "copy_edge", "sum", "max",
           "u_add_v", "u_mul_v", "u_dot_v", "copy_u", "copy_v",
           "BuiltinFunction", "BundledFunction"]


class BuiltinFunction(object):
    """Base class of builtins."""

    @property
    def name(self):
        raise NotImplementedError


class BundledFunction(object):
    """Calls several functions and merges their result dicts (function/base.py:13-38)."""

    def __init__(self, fn_list):
        self.fn_list = fn_list

    def __call__(self, *args, **kwargs):
        out = {}
        for fn in self.fn_list:
            out.update(fn(*args, **kwargs))
        return out

    @property
    def name(self):
        return "bundled"


class MessageFunction(BuiltinFunction):
    """Base class of builtin message functions."""

    def __call__(self, edges):
        raise NotImplementedError

    def is_spmv_supported(self, g):
        raise NotImplementedError

    @property
    def use_edge_feature(self):
        raise NotImplementedError


def _weight_width(src_shape, edge_shape):
    """How a src_mul_edge weight of `edge_shape[1:]` broadcasts over a source feature of
    `src_shape[1:]` after the reference's rank alignment (append 1s, message.py:82-88).
    Returns w_width such that flat feature f uses weight f // (feat // w_width), or None when
    the broadcast is not of that trailing-group form (then the UDF path is used)."""
    s = tuple(src_shape[1:])
    e = tuple(edge_shape[1:])
    rank = builtins.max(len(s), len(e))
    s = s + (1,) * (rank - len(s))
    e = e + (1,) * (rank - len(e))
    # leading dims must match exactly, then the weight must be all ones
    k = 0
    while k < rank and e[k] == s[k]:
        k += 1
    if any(d != 1 for d in e[k:]):
        return None
    width = 1
    for d in e[:k]:
        width *= d
    return width



class SrcMulEdgeMessageFunction(MessageFunction):
    """message = src[src_field] * edge[edge_field] with rank alignment."""

    def __init__(self, mul_op, src_field, edge_field, out_field):
        self.mul_op = mul_op
        self.src_field = src_field
        self.edge_field = edge_field
        self.out_field = out_field

    def is_spmv_supported(self, g):
        return self.kernel_form(g) is not None

    def kernel_form(self, g):
        """(w_width) if the pair lowers to the SRC_MUL_EDGE kernel, else None."""
        x = g.ndata[self.src_field]
        w = g.edata[self.edge_field]
        return _weight_width(x.shape, w.shape)

    def __call__(self, edges):
        sdata = edges.src[self.src_field]
        edata = edges.data[self.edge_field]
        rank = builtins.max(sdata.dim(), edata.dim())
        sdata = sdata.reshape(tuple(sdata.shape) + (1,) * (rank - sdata.dim()))
        edata = edata.reshape(tuple(edata.shape) + (1,) * (rank - edata.dim()))
        return {self.out_field: self.mul_op(sdata, edata)}

    @property
    def name(self):
        return "src_mul_edge"

    @property
    def use_edge_feature(self):
        return True


class CopySrcMessageFunction(MessageFunction):
    """message = src[src_field]."""

    def __init__(self, src_field, out_field):
        self.src_field = src_field
        self.out_field = out_field

    def is_spmv_supported(self, g):
        return True

    def __call__(self, edges):
        return {self.out_field: edges.src[self.src_field]}

    @property
    def name(self):
        return "copy_src"

    @property
    def use_edge_feature(self):
        return False


class CopyEdgeMessageFunction(MessageFunction):
    """message = edge[edge_field]."""

    def __init__(self, edge_field=None, out_field=None):
        self.edge_field = edge_field
        self.out_field = out_field

    def is_spmv_supported(self, g):
        # the reference hard-codes False (message.py:156-170); the COPY_EDGE kernel covers it
        return True

    def __call__(self, edges):
        return {self.out_field: edges.data[self.edge_field]}

    @property
    def name(self):
        return "copy_edge"

    @property
    def use_edge_feature(self):
        return True


def src_mul_edge(src, edge, out):
    """Builtin message: edges.src[src] * edges.data[edge] -> out."""
    return SrcMulEdgeMessageFunction(operator.mul, src, edge, out)


def copy_src(src, out):
    """Builtin message: edges.src[src] -> out."""
    return CopySrcMessageFunction(src, out)


def copy_edge(edge, out):
    """Builtin message: edges.data[edge] -> out."""
    return CopyEdgeMessageFunction(edge, out)


class ReduceFunction(BuiltinFunction):
    """Base class of builtin reducers."""

    def __call__(self, nodes):
        raise NotImplementedError

    def is_spmv_supported(self):
        raise NotImplementedError


def _torch_sum(x, dim):
    return torch.sum(x, dim)


def _torch_max(x, dim):
    # backend F.max drops the argmax (backend/pytorch/tensor.py:95-97)
    return torch.max(x, dim)[0]


class SimpleReduceFunction(ReduceFunction):
    """Reduce one message field into one node field."""

    def __init__(self, name, reduce_op, msg_field, out_field):
        self._name = name
        self.reduce_op = reduce_op
        self.msg_field = msg_field
        self.out_field = out_field

    def is_spmv_supported(self):
        # the reference supports only "sum" (reducer.py:40-43)
        return self._name in ("sum", "max")

    def __call__(self, nodes):
        return {self.out_field: self.reduce_op(nodes.mailbox[self.msg_field], 1)}

    @property
    def name(self):
        return self._name


def sum(msg, out):  # pylint: disable=redefined-builtin
    """Builtin reducer: sum of mailbox[msg] -> out."""
    return SimpleReduceFunction("sum", _torch_sum, msg, out)


def max(msg, out):  # pylint: disable=redefined-builtin
    """Builtin reducer: max of mailbox[msg] -> out."""
    return SimpleReduceFunction("max", _torch_max, msg, out)


# ---- builtin edge functions -------------------------------------------------
# The reference (0.1.3) has no builtin for apply_edges: the GAT example computes
# a1[src] + a2[dst] in a Python edge UDF over three materialised E-row gathers
# (examples/pytorch/gat/train.py:90-97 through runtime/scheduler.py:234-277).  These builtins name the
# same per-edge expressions so that apply_edges lowers them to ONE launch of the SDDMM kernel
# (dglb200_sddmm_csr); called as a UDF they compute exactly what the UDF would.

class EdgeFunction(BuiltinFunction):
    """u-op-v over (edges.src[lhs], edges.dst[rhs]) -> edges.data[out]."""

    def __init__(self, name, lhs_field, rhs_field, out_field):
        self._name = name
        self.lhs_field = lhs_field
        self.rhs_field = rhs_field
        self.out_field = out_field

    @property
    def name(self):
        return self._name

    def __call__(self, edges):
        if self._name == "copy_u":
            return {self.out_field: edges.src[self.lhs_field]}
        if self._name == "copy_v":
            return {self.out_field: edges.dst[self.rhs_field]}
        u, v = edges.src[self.lhs_field], edges.dst[self.rhs_field]
        if self._name == "u_add_v":
            return {self.out_field: u + v}
        if self._name == "u_mul_v":
            return {self.out_field: u * v}
        return {self.out_field: (u * v).sum(-1, keepdim=True)}      # u_dot_v over the last dim


def u_add_v(lhs, rhs, out):
    """Builtin edge function: edges.src[lhs] + edges.dst[rhs] -> out."""
    return EdgeFunction("u_add_v", lhs, rhs, out)


def u_mul_v(lhs, rhs, out):
    """Builtin edge function: edges.src[lhs] * edges.dst[rhs] -> out."""
    return EdgeFunction("u_mul_v", lhs, rhs, out)


def u_dot_v(lhs, rhs, out):
    """Builtin edge function: <edges.src[lhs], edges.dst[rhs]> over the last dim -> out (..., 1)."""
    return EdgeFunction("u_dot_v", lhs, rhs, out)


def copy_u(lhs, out):
    """Builtin edge function: edges.src[lhs] -> out."""
    return EdgeFunction("copy_u", lhs, None, out)


def copy_v(rhs, out):
    """Builtin edge function: edges.dst[rhs] -> out."""
    return EdgeFunction("copy_v", None, rhs, out)
