"""Thin tensor-level wrappers over the C ABI (include/dgl_b200.h).

This is the layer that plays `F.spmm` / `F.sparse_matrix` of the reference's backend shim
(python/dgl/backend/pytorch/tensor.py:34-54,138-146): torch supplies device memory and the
current CUDA stream, every computation is a libdgl_b200.so kernel.  No function here has a
CPU or eager-PyTorch fallback; a non-CUDA tensor raises DGLError.
"""
import ctypes

import torch

from . import _lib
from .base import DGLError

MSG_COPY_SRC, MSG_SRC_MUL_EDGE, MSG_COPY_EDGE = 0, 1, 2
RED_SUM, RED_MAX = 0, 1
F32, BF16 = 0, 1
SDDMM_DOT, SDDMM_ADD, SDDMM_MUL, SDDMM_COPY_U, SDDMM_COPY_V = 0, 1, 2, 3, 4
SOFTMAX_STABLE, SOFTMAX_GAT_EXAMPLE = 0, 1

DEFAULT_CHUNK = 512
_BAD_FLAGS = {}      # per device: int32 flag the gather / select kernels raise on an out-of-range id


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    if t is None:
        return ctypes.c_void_p(0)
    return ctypes.c_void_p(t.data_ptr())


def _need_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise DGLError("dgl_b200 ops need CUDA tensors (got %s); there is no CPU path" % t.device)


def _dtype_code(t):
    if t.dtype == torch.float32:
        return F32
    if t.dtype == torch.bfloat16:
        return BF16
    raise DGLError("unsupported feature dtype %s (float32 / bfloat16 only)" % t.dtype)


def _workspace(nbytes, device):
    return torch.empty(max(int(nbytes), 1), dtype=torch.uint8, device=device)


class CSR(object):
    """Engine-side CSR: int64 indptr[n_rows+1], int32 indices[nnz], int32 eids[nnz]."""

    def __init__(self, n_rows, n_cols, indptr, indices, eids):
        self.n_rows = int(n_rows)
        self.n_cols = int(n_cols)
        self.indptr = indptr
        self.indices = indices
        self.eids = eids
        self.nnz = int(indices.shape[0])
        self._plans = {}

    @property
    def device(self):
        return self.indptr.device

    def plan(self, chunk=DEFAULT_CHUNK):
        """Long-row split table for this CSR (cached per chunk)."""
        if chunk not in self._plans:
            self._plans[chunk] = SpmmPlan(self, chunk)
        return self._plans[chunk]


def csr_build(row, col, n_rows, n_cols, eid=None, order=0):
    """dglb200_csr_build: CSR keyed by `row` of the COO (row, col[, eid]) (int64, CUDA)."""
    _need_cuda(row, col, eid)
    lib = _lib.get_lib()
    nnz = int(row.shape[0])
    if row.dtype != torch.int64 or col.dtype != torch.int64 or (eid is not None and eid.dtype != torch.int64):
        raise DGLError("csr_build: ids must be int64 (c_api_common.h:35-38 IsValidIdArray)")
    row = row.contiguous()
    col = col.contiguous()
    eid = eid.contiguous() if eid is not None else None
    dev = row.device
    indptr = torch.empty(n_rows + 1, dtype=torch.int64, device=dev)
    indices = torch.empty(nnz, dtype=torch.int32, device=dev)
    eids = torch.empty(nnz, dtype=torch.int32, device=dev)
    nbytes = ctypes.c_size_t(0)
    _lib.check_call(lib.dglb200_csr_build_workspace_bytes(n_rows, nnz, ctypes.byref(nbytes)))
    ws = _workspace(nbytes.value, dev)
    with torch.cuda.device(dev):
        _lib.check_call(lib.dglb200_csr_build(n_rows, n_cols, nnz, _ptr(row), _ptr(col), _ptr(eid),
                                              order, _ptr(indptr), _ptr(indices), _ptr(eids),
                                              _ptr(ws), nbytes.value, _stream()))
    return CSR(n_rows, n_cols, indptr, indices, eids)


def csr_row_block(csr, r0, r1):
    """Rows [r0, r1) of a CSR as a CSR over the same columns (views of indices / eids; the row
    pointer is rebased).  This is the 1-D destination-row partition of SURVEY.md 8e."""
    lo = int(csr.indptr[r0])
    hi = int(csr.indptr[r1])
    indptr = (csr.indptr[r0:r1 + 1] - lo).contiguous()
    return CSR(r1 - r0, csr.n_cols, indptr, csr.indices[lo:hi], csr.eids[lo:hi])


def widen(t):
    """int32 engine array -> int64 (the reference's id dtype)."""
    _need_cuda(t)
    out = torch.empty(t.shape, dtype=torch.int64, device=t.device)
    with torch.cuda.device(t.device):
        _lib.check_call(_lib.get_lib().dglb200_widen_i32_to_i64(t.numel(), _ptr(t), _ptr(out), _stream()))
    return out


def degrees(csr):
    out = torch.empty(csr.n_rows, dtype=torch.int64, device=csr.device)
    with torch.cuda.device(csr.device):
        _lib.check_call(_lib.get_lib().dglb200_degrees_from_indptr(csr.n_rows, _ptr(csr.indptr), _ptr(out), _stream()))
    return out


def csr_row_ids(csr):
    out = torch.empty(csr.nnz, dtype=torch.int64, device=csr.device)
    with torch.cuda.device(csr.device):
        _lib.check_call(_lib.get_lib().dglb200_csr_row_ids(csr.n_rows, csr.nnz, _ptr(csr.indptr), _ptr(out), _stream()))
    return out


def csr_select(csr, rows, want=("which", "slot")):
    """Edge lists of the CSR rows `rows` (int64 CUDA, in the order given): dglb200_csr_select_offsets +
    dglb200_csr_select_expand.  `want` names the outputs, returned in that order: "which" (position in
    `rows`), "slot" (CSR slot), "row" (= rows[which]), "col" (column id), "eid" (edge id); all int64."""
    lib = _lib.get_lib()
    _need_cuda(rows, csr.indptr)
    rows = rows.contiguous()
    if rows.dtype != torch.int64:
        raise DGLError("csr_select: ids must be int64")
    dev = csr.device
    n_sel = int(rows.shape[0])
    nbytes = ctypes.c_size_t(0)
    _lib.check_call(lib.dglb200_csr_select_workspace_bytes(n_sel, ctypes.byref(nbytes)))
    ws = _workspace(nbytes.value, dev)
    offsets = torch.empty(n_sel + 1, dtype=torch.int64, device=dev)
    total = ctypes.c_int64(0)
    with torch.cuda.device(dev):
        _lib.check_call(lib.dglb200_csr_select_offsets(n_sel, _ptr(rows), csr.n_rows, _ptr(csr.indptr),
                                                       _ptr(offsets), ctypes.byref(total), _ptr(ws),
                                                       nbytes.value, _stream()))
    total = int(total.value)
    outs = {k: (torch.empty(total, dtype=torch.int64, device=dev) if k in want else None)
            for k in ("which", "slot", "row", "col", "eid")}
    with torch.cuda.device(dev):
        _lib.check_call(lib.dglb200_csr_select_expand(
            n_sel, total, _ptr(rows), _ptr(offsets), _ptr(csr.indptr), _ptr(csr.indices), _ptr(csr.eids),
            _ptr(outs["which"]), _ptr(outs["slot"]), _ptr(outs["row"]), _ptr(outs["col"]), _ptr(outs["eid"]),
            _stream()))
    return tuple(outs[k] for k in want)


def neighbor_sample(csr, rows, fanout, seed, want=("row", "col", "eid")):
    """One hop of uniform neighbour sampling: min(degree, fanout) distinct slots of every CSR row in `rows`
    (int64 CUDA, each row at most once per call), ascending slot order per row, rows in the order given:
    dglb200_neighbor_sample_offsets + dglb200_neighbor_sample_fill (without the device-side frontier / discovery
    arguments, which dglb200_sample_subgraph drives itself).  Returns (offsets, outputs named by `want` out of
    "row", "col", "eid", "slot"; all int64)."""
    lib = _lib.get_lib()
    _need_cuda(rows, csr.indptr)
    rows = rows.contiguous()
    if rows.dtype != torch.int64:
        raise DGLError("neighbor_sample: ids must be int64")
    dev = csr.device
    n_sel = int(rows.shape[0])
    nbytes = ctypes.c_size_t(0)
    _lib.check_call(lib.dglb200_csr_select_workspace_bytes(n_sel, ctypes.byref(nbytes)))
    ws = _workspace(nbytes.value, dev)
    offsets = torch.empty(n_sel + 1, dtype=torch.int64, device=dev)
    total = ctypes.c_int64(0)
    with torch.cuda.device(dev):
        _lib.check_call(lib.dglb200_neighbor_sample_offsets(n_sel, _ptr(rows), None, csr.n_rows, _ptr(csr.indptr),
                                                            int(fanout), _ptr(offsets), ctypes.byref(total), None,
                                                            _ptr(ws), nbytes.value, _stream()))
    total = int(total.value)
    outs = {k: (torch.empty(total, dtype=torch.int64, device=dev) if k in want else None)
            for k in ("row", "col", "eid", "slot")}
    with torch.cuda.device(dev):
        _lib.check_call(lib.dglb200_neighbor_sample_fill(
            n_sel, _ptr(rows), None, _ptr(offsets), _ptr(csr.indptr), _ptr(csr.indices), _ptr(csr.eids),
            int(fanout), int(seed) & 0xFFFFFFFFFFFFFFFF, _ptr(outs["row"]), _ptr(outs["col"]), _ptr(outs["eid"]),
            _ptr(outs["slot"]), None, None, 0, None, None, _stream()))
    return (offsets,) + tuple(outs[k] for k in want)


class SampleState(object):
    """Persistent discovery state of the sampler for one parent graph (dglb200_sample_subgraph):
    layer int32[n] = -1, bitmap uint32[ceil(n / 32)] = 0, word-prefix scratch.  One sampling call at a time."""

    def __init__(self, n_rows, device):
        self.n_rows = int(n_rows)
        words = (self.n_rows + 31) // 32
        self.layer = torch.full((self.n_rows,), -1, dtype=torch.int32, device=device)
        self.bitmap = torch.zeros(words, dtype=torch.int32, device=device)
        self.word_prefix = torch.empty(words + 1, dtype=torch.int32, device=device)
        self.ws_bytes = 0          # workspace size that carried the largest call so far


def sample_subgraph(csr, state, seeds, fanout, num_hops, seed):
    """dglb200_sample_subgraph: the whole multi-hop call.  Returns (nodes, layers, sub_indptr, sub_indices int32,
    sub_rows, induced_edges) - views into one workspace tensor that the call's results live in.  The workspace
    is sized from the largest call seen on this state; a call that needs more is retried once it says how much."""
    lib = _lib.get_lib()
    _need_cuda(seeds, csr.indptr, state.layer)
    dev = csr.device
    seeds = seeds.contiguous()
    out = _lib.SampledSubgraphStruct()
    guess = max(state.ws_bytes, 1 << 16)
    for _ in range(2 * int(num_hops) + 4):      # at most two shortfalls per hop (offsets, picks) + the outputs
        ws = _workspace(guess, dev)
        with torch.cuda.device(dev):
            rc = lib.dglb200_sample_subgraph(
                csr.n_rows, _ptr(csr.indptr), _ptr(csr.indices), _ptr(csr.eids), int(seeds.shape[0]), _ptr(seeds),
                int(fanout), int(num_hops), int(seed) & 0xFFFFFFFFFFFFFFFF, _ptr(state.layer), _ptr(state.bitmap),
                _ptr(state.word_prefix), _ptr(ws), guess, ctypes.byref(out), _stream())
        if rc == 0:
            break
        if out.needed_bytes <= guess:
            raise DGLError(_lib.last_error())
        guess = int(out.needed_bytes)
    else:
        raise DGLError(_lib.last_error())
    state.ws_bytes = max(state.ws_bytes, int(out.needed_bytes) * 5 // 4)
    base = ws.data_ptr()

    def view(ptr, count, dtype):
        if count == 0 or not ptr:
            return torch.empty(0, dtype=dtype, device=dev)
        off = ptr - base
        return ws[off:off + count * dtype.itemsize].view(dtype)
    n, e = int(out.n_nodes), int(out.n_edges)
    return (view(out.nodes, n, torch.int64), view(out.layers, n, torch.int64), view(out.sub_indptr, n + 1, torch.int64),
            view(out.sub_indices, e, torch.int32), view(out.sub_rows, e, torch.int64),
            view(out.induced_edges, e, torch.int64))


def degrees_of(csr, rows):
    """Degrees of the CSR rows `rows` (int64 CUDA): dglb200_degrees_of."""
    _need_cuda(rows, csr.indptr)
    rows = rows.contiguous()
    out = torch.empty(rows.shape[0], dtype=torch.int64, device=csr.device)
    bad = _BAD_FLAGS.get(csr.device)
    if bad is None:
        bad = _BAD_FLAGS[csr.device] = torch.zeros(1, dtype=torch.int32, device=csr.device)
    with torch.cuda.device(csr.device):
        _lib.check_call(_lib.get_lib().dglb200_degrees_of(int(rows.shape[0]), _ptr(rows), csr.n_rows,
                                                          _ptr(csr.indptr), _ptr(out), _ptr(bad), _stream()))
    return out


class SpmmPlan(object):
    """Host handle of dglb200_spmm_plan_t; owns the device tables and the partial buffers."""

    def __init__(self, csr, chunk):
        lib = _lib.get_lib()
        self.chunk = int(chunk)
        self.csr = csr
        dev = csr.device
        nbytes = ctypes.c_size_t(0)
        _lib.check_call(lib.dglb200_spmm_plan_workspace_bytes(csr.n_rows, ctypes.byref(nbytes)))
        ws = _workspace(nbytes.value, dev)
        # upper bound on long rows: every long row has > chunk edges
        cap = csr.nnz // self.chunk + 1
        self.long_rows = torch.empty(cap, dtype=torch.int32, device=dev)
        self.chunk_ptr = torch.empty(cap + 1, dtype=torch.int64, device=dev)
        counts = (ctypes.c_int64 * 2)()
        with torch.cuda.device(dev):
            _lib.check_call(lib.dglb200_spmm_plan(csr.n_rows, _ptr(csr.indptr), self.chunk, cap,
                                                  _ptr(self.long_rows), _ptr(self.chunk_ptr), counts,
                                                  _ptr(ws), nbytes.value, _stream()))
        self.n_long_rows = int(counts[0])
        self.n_chunks = int(counts[1])

    def struct(self, feat, want_arg):
        """Fill a dglb200_spmm_plan_t for a call with `feat` features (None if no long rows).  The fp32
        scratch of the long-row slices is allocated PER CALL from torch's stream-aware caching allocator
        (a cached block, no cudaMalloc in steady state): two SpMMs over the same CSR issued on different
        streams (the exchange pipelines of dist.py, user streams) no longer share one buffer (round-1 advisor
        finding).  The struct keeps the tensors alive until the caller has enqueued the kernels; the
        allocator then reuses the block in stream order."""
        if self.n_long_rows == 0:
            return None
        need = self.n_chunks * feat
        dev = self.csr.device
        partial = torch.empty(need, dtype=torch.float32, device=dev)
        partial_arg = torch.empty(need, dtype=torch.int32, device=dev) if want_arg else None
        st = _lib.SpmmPlanStruct()
        st.chunk = self.chunk
        st.n_long_rows = self.n_long_rows
        st.n_chunks = self.n_chunks
        st.long_rows = self.long_rows.data_ptr()
        st.chunk_ptr = self.chunk_ptr.data_ptr()
        st.partial = partial.data_ptr()
        st.partial_arg = partial_arg.data_ptr() if want_arg else None
        st.partial_bytes = need * 4
        st._keep = (partial, partial_arg)
        return st


def _flatten(t, n):
    """(n, d1, d2, ...) -> (n, prod d) as SPMVExecutor.run does (ir/executor.py:457-472)."""
    if t.dim() == 1:
        return t.reshape(n, 1)
    return t.reshape(n, -1)


def _pad_rows(x, multiple):
    """Row-major copy of x whose leading dimension is rounded up to `multiple` elements (zero pad);
    returns the (rows, feat) view into it."""
    rows, feat = x.shape
    ld = (feat + multiple - 1) // multiple * multiple
    buf = torch.zeros((rows, ld), dtype=x.dtype, device=x.device)
    buf[:, :feat].copy_(x)
    return buf[:, :feat]


SLOT_ORDER_MIN_NNZ = 1 << 22   # below this the per-slot w[eid] gather stays in L2 and costs nothing


def edge_data_to_slots(csr, w):
    """w (E,) or (E, k) in edge-id order -> the same rows in the slot order of `csr`
    (dglb200_edge_data_to_slots).  Identity when the CSR carries no edge ids."""
    lib = _lib.get_lib()
    _need_cuda(w, csr.indptr)
    w = w.contiguous()
    if csr.eids is None:
        return w
    width = w.numel() // max(w.shape[0], 1)
    out = torch.empty((csr.nnz,) + tuple(w.shape[1:]), dtype=w.dtype, device=w.device)
    with torch.cuda.device(w.device):
        _lib.check_call(lib.dglb200_edge_data_to_slots(
            csr.nnz, _ptr(csr.eids), w.shape[0], width * w.element_size(), _ptr(w), _ptr(out), None,
            _stream()))
    return out


def _slot_ordered_weights(csr, w, w_width, feat, need_eids):
    """(w, eids) to hand to the kernel: on large graphs narrow edge weights (a scalar or one per
    head) are permuted to slot order once, so that every feature slab reads them coalesced."""
    if (w is None or need_eids or csr.eids is None or csr.nnz < SLOT_ORDER_MIN_NNZ
            or w_width >= feat or w_width * w.element_size() > 64):
        return w, csr.eids
    return edge_data_to_slots(csr, w.reshape(w.shape[0], -1)), None


def spmm(csr, msg_op, reduce_op, x=None, w=None, w_width=1, want_arg=False, chunk=DEFAULT_CHUNK,
         out=None, accumulate=False, x_halo=None, halo_first=0, row_scale=None, bias=None, relu=False):
    """y[r] = reduce over row r of msg(x[indices], w[eids]); 2-D (rows, feat) tensors.

    Returns (y, arg) where arg is int32 (rows, feat) for RED_MAX with want_arg, else None.
    accumulate: y (= out) += result (SUM).  x_halo / halo_first: column ids >= halo_first are rows of the
    bfloat16 tensor x_halo (fp32 copy_src -> sum only; dglb200_spmm_csr_ext).
    row_scale (rows,) / bias (feat,) fp32 / relu: the SUM epilogue y = relu(row_scale[r] * sum + bias[f]).
    """
    lib = _lib.get_lib()
    ref = x if x is not None else w
    _need_cuda(ref, x, w, csr.indptr)
    if msg_op == MSG_COPY_EDGE:
        feat = int(w.shape[1])
        w_width = feat
    else:
        feat = int(x.shape[1])
    dt = _dtype_code(ref)

    def rows_ok(t):    # row-major with unit inner stride (a column block of a wider matrix is fine)
        return t.dim() == 2 and t.stride(1) == 1 and t.stride(0) >= t.shape[1]
    if x is not None:
        if not rows_ok(x):
            x = x.contiguous()
        if x_halo is None and x.shape[0] != csr.n_cols:
            raise DGLError("spmm: x has %d rows, CSR has %d columns" % (x.shape[0], csr.n_cols))
        if x_halo is not None and (x_halo.dtype != torch.bfloat16 or not x_halo.is_contiguous()
                                   or x_halo.shape[1] != x.shape[1] or halo_first > x.shape[0]):
            raise DGLError("spmm: x_halo must be a contiguous bfloat16 (rows, feat) tensor and halo_first <= len(x)")
    if w is not None:
        w = w.contiguous()
        if w.dtype != ref.dtype:
            raise DGLError("spmm: edge data dtype %s != feature dtype %s" % (w.dtype, ref.dtype))
    # ragged widths (602 = Reddit's input features): pad the gathered rows to a 16-byte multiple so
    # the kernel keeps its 128-bit loads; one n x feat copy against nnz x feat gathered elements
    vec = 16 // ref.element_size()
    if (x is not None and feat % vec != 0 and feat > 2 * vec and csr.nnz >= 4 * csr.n_cols
            and (msg_op == MSG_COPY_SRC or (msg_op == MSG_SRC_MUL_EDGE and w_width == 1))
            and (x.stride(0) % vec != 0 or x.data_ptr() % 16 != 0)):
        x = _pad_rows(x, vec)
    if out is not None and not rows_ok(out):
        raise DGLError("spmm: `out` must be row-major with unit inner stride")
    y = out if out is not None else torch.empty((csr.n_rows, feat), dtype=ref.dtype, device=ref.device)
    arg = None
    if reduce_op == RED_MAX and want_arg:
        arg = torch.empty((csr.n_rows, feat), dtype=torch.int32, device=ref.device)
    plan = csr.plan(chunk).struct(feat, arg is not None) if chunk else None
    plan_ref = ctypes.byref(plan) if plan is not None else None
    x_ld = int(x.stride(0)) if (x is not None and x.shape[0] > 1) else feat
    y_ld = int(y.stride(0)) if y.shape[0] > 1 else feat
    eids = csr.eids
    if msg_op == MSG_SRC_MUL_EDGE:
        w, eids = _slot_ordered_weights(csr, w, w_width, feat, need_eids=arg is not None)
    ext_ref = None
    if accumulate or x_halo is not None or row_scale is not None or bias is not None or relu:
        for t_, n_, name in ((row_scale, csr.n_rows, "row_scale"), (bias, feat, "bias")):
            if t_ is not None and (t_.dtype != torch.float32 or not t_.is_contiguous() or t_.numel() != n_):
                raise DGLError("spmm: %s must be a contiguous float32 tensor of %d elements" % (name, n_))
        if accumulate and out is None:
            raise DGLError("spmm: accumulate needs `out`")
        ext = _lib.SpmmExtStruct()
        ext.accumulate = 1 if accumulate else 0
        ext.x_halo = x_halo.data_ptr() if x_halo is not None else None
        ext.halo_first = int(halo_first)
        ext.row_scale = row_scale.data_ptr() if row_scale is not None else None
        ext.bias = bias.data_ptr() if bias is not None else None
        ext.relu = 1 if relu else 0
        ext_ref = ctypes.byref(ext)
    with torch.cuda.device(ref.device):
        _lib.check_call(lib.dglb200_spmm_csr_ext(
            csr.n_rows, csr.n_cols, csr.nnz, _ptr(csr.indptr), _ptr(csr.indices), _ptr(eids),
            msg_op, reduce_op, dt, _ptr(x), max(x_ld, feat), feat, _ptr(w), w_width, _ptr(y),
            max(y_ld, feat), _ptr(arg), plan_ref, ext_ref, _stream()))
    return y, arg


def spmm_backward_x(csr_t, dy, w=None, w_width=1, arg=None, chunk=DEFAULT_CHUNK):
    """dx over the CSR keyed by source (see dglb200_spmm_csr_backward_x)."""
    lib = _lib.get_lib()
    _need_cuda(dy, w, arg)
    feat = int(dy.shape[1])
    if arg is None and feat % (16 // dy.element_size()) != 0:
        # the unmasked gradient IS the forward op over the transposed CSR; ragged widths take the
        # row-padding route of spmm()
        msg = MSG_COPY_SRC if w is None else MSG_SRC_MUL_EDGE
        return spmm(csr_t, msg, RED_SUM, x=dy, w=w, w_width=w_width, chunk=chunk)[0]
    dy = dy.contiguous()
    dx = torch.empty((csr_t.n_rows, feat), dtype=dy.dtype, device=dy.device)
    plan = csr_t.plan(chunk).struct(feat, False) if chunk else None
    plan_ref = ctypes.byref(plan) if plan is not None else None
    eids = csr_t.eids
    if w is not None:
        w = w.contiguous()
        w, eids = _slot_ordered_weights(csr_t, w, w_width, feat, need_eids=arg is not None)
    with torch.cuda.device(dy.device):
        _lib.check_call(lib.dglb200_spmm_csr_backward_x(
            csr_t.n_rows, csr_t.n_cols, csr_t.nnz, _ptr(csr_t.indptr), _ptr(csr_t.indices), _ptr(eids),
            _dtype_code(dy), _ptr(dy), feat, _ptr(w), w_width, _ptr(arg), _ptr(dx), plan_ref,
            _stream()))
    return dx


def sddmm(csr, op, u, v, heads, dim):
    """Per-edge u-op-v over the in-edge CSR (dglb200_sddmm_csr).  u: (n_cols, heads*dim),
    v: (n_rows, heads*dim).  DOT -> (nnz, heads); ADD / MUL -> (nnz, heads*dim); by edge id."""
    lib = _lib.get_lib()
    _need_cuda(u, v)
    u = u.contiguous() if u is not None else None
    v = v.contiguous() if v is not None else None
    ref = u if u is not None else v
    if u is not None and v is not None and u.dtype != v.dtype:
        raise DGLError("sddmm: dtype mismatch %s vs %s" % (u.dtype, v.dtype))
    width = heads if op == SDDMM_DOT else heads * dim
    out = torch.empty((csr.nnz, width), dtype=ref.dtype, device=ref.device)
    with torch.cuda.device(ref.device):
        _lib.check_call(lib.dglb200_sddmm_csr(csr.n_rows, csr.nnz, _ptr(csr.indptr), _ptr(csr.indices),
                                              _ptr(csr.eids), op, _dtype_code(ref), _ptr(u), _ptr(v),
                                              heads, dim, _ptr(out), _stream()))
    return out


def edge_softmax(csr, score):
    """Stable per-destination softmax of (nnz, heads) fp32 scores indexed by edge id."""
    lib = _lib.get_lib()
    _need_cuda(score)
    score = score.contiguous().float()
    heads = int(score.shape[1])
    alpha = torch.zeros_like(score)
    with torch.cuda.device(score.device):
        _lib.check_call(lib.dglb200_edge_softmax_csr(csr.n_rows, csr.nnz, _ptr(csr.indptr), _ptr(csr.eids),
                                                     heads, _ptr(score), _ptr(alpha), _stream()))
    return alpha


def edge_softmax_backward(csr, alpha, dalpha):
    lib = _lib.get_lib()
    _need_cuda(alpha, dalpha)
    alpha = alpha.contiguous()
    dalpha = dalpha.contiguous()
    heads = int(alpha.shape[1])
    dscore = torch.zeros_like(alpha)
    with torch.cuda.device(alpha.device):
        _lib.check_call(lib.dglb200_edge_softmax_backward_csr(csr.n_rows, csr.nnz, _ptr(csr.indptr),
                                                              _ptr(csr.eids), heads, _ptr(alpha),
                                                              _ptr(dalpha), _ptr(dscore), _stream()))
    return dscore


def gat_aggregate(csr, a_src, a_dst, ft, negative_slope, flavour=SOFTMAX_GAT_EXAMPLE, want_edge=True):
    """Fused GAT edge path (dglb200_gat_aggregate_csr).  a_src: (n_cols, H), a_dst: (n_rows, H),
    ft: (n_cols, H, D) fp32.  Returns (out (n_rows, H, D), a (nnz, H) or None, z (n_rows, H))."""
    lib = _lib.get_lib()
    _need_cuda(a_src, a_dst, ft)
    heads, dim = int(ft.shape[1]), int(ft.shape[2])
    a_src = a_src.reshape(-1, heads).contiguous().float()
    a_dst = a_dst.reshape(-1, heads).contiguous().float()
    ft = ft.contiguous().float()
    out = torch.empty((csr.n_rows, heads, dim), dtype=torch.float32, device=ft.device)
    a = torch.zeros((csr.nnz, heads), dtype=torch.float32, device=ft.device) if want_edge else None
    z = torch.empty((csr.n_rows, heads), dtype=torch.float32, device=ft.device)
    with torch.cuda.device(ft.device):
        _lib.check_call(lib.dglb200_gat_aggregate_csr(
            csr.n_rows, csr.nnz, _ptr(csr.indptr), _ptr(csr.indices), _ptr(csr.eids), _ptr(a_src),
            _ptr(a_dst), _ptr(ft), heads, dim, float(negative_slope), flavour, _ptr(out), _ptr(a),
            _ptr(z), _stream()))
    return out, a, z


def degree_bucketing(msg_ids, vids, recv_ids, id_bound):
    """dglb200_degree_bucketing: canonical (degs, nids, nid_section, mids, mid_section), int64 CUDA."""
    lib = _lib.get_lib()
    _need_cuda(msg_ids, vids, recv_ids)
    dev = vids.device
    m, r = int(vids.shape[0]), int(recv_ids.shape[0])
    msg_ids = msg_ids.contiguous()
    vids = vids.contiguous()
    recv_ids = recv_ids.contiguous()
    cap = int(lib.dglb200_bucket_capacity(m))
    i64 = dict(dtype=torch.int64, device=dev)
    degs = torch.empty(cap, **i64)
    nsec = torch.empty(cap, **i64)
    msec = torch.empty(cap, **i64)
    nids = torch.empty(m + r, **i64)
    mids = torch.empty(m, **i64)
    nbytes = ctypes.c_size_t(0)
    _lib.check_call(lib.dglb200_degree_bucketing_workspace_bytes(m, r, ctypes.byref(nbytes)))
    ws = _workspace(nbytes.value, dev)
    counts = (ctypes.c_int64 * 3)()
    with torch.cuda.device(dev):
        _lib.check_call(lib.dglb200_degree_bucketing(
            m, r, int(id_bound), _ptr(msg_ids), _ptr(vids), _ptr(recv_ids), _ptr(degs), _ptr(nids),
            _ptr(nsec), _ptr(mids), _ptr(msec), counts, _ptr(ws), nbytes.value, _stream()))
    b, k, bnz = int(counts[0]), int(counts[1]), int(counts[2])
    return degs[:b], nids[:k], nsec[:b], mids, msec[:bnz]


def unique_sorted(ids, id_bound, want_inverse=True):
    """dglb200_unique_sorted: (sorted unique ids, inverse positions)."""
    lib = _lib.get_lib()
    _need_cuda(ids)
    ids = ids.contiguous()
    n = int(ids.shape[0])
    dev = ids.device
    uniq = torch.empty(n, dtype=torch.int64, device=dev)
    inv = torch.empty(n, dtype=torch.int64, device=dev) if want_inverse else None
    nbytes = ctypes.c_size_t(0)
    _lib.check_call(lib.dglb200_unique_workspace_bytes(n, ctypes.byref(nbytes)))
    ws = _workspace(nbytes.value, dev)
    counts = (ctypes.c_int64 * 1)()
    with torch.cuda.device(dev):
        _lib.check_call(lib.dglb200_unique_sorted(n, int(id_bound), _ptr(ids), _ptr(uniq), _ptr(inv),
                                                  counts, _ptr(ws), nbytes.value, _stream()))
    return uniq[:int(counts[0])], inv


def gather_rows(x, idx, out=None, max_ctas=0):
    """out[i] = x[idx[i]] (idx int64 CUDA), or out[i] = x[i] when idx is None (strided block
    copy).  x / out may be column blocks of wider row-major matrices.  The kernel runs on the
    device (and current stream) of `out`; `x` may alias a PEER GPU's memory mapped into this
    process (dist.share_with_peers): the rows are then pulled over NVLink by the gathering GPU.
    max_ctas > 0 bounds the grid so that a concurrent SpMM keeps the SMs."""
    lib = _lib.get_lib()
    _need_cuda(x, idx, out)
    tail = tuple(x.shape[1:])
    if x.dim() == 1:
        x = x.reshape(-1, 1)
    if x.dim() > 2:
        if not x.is_contiguous():
            raise DGLError("gather_rows: N-d x must be contiguous")
        x2 = x.reshape(x.shape[0], -1)
    else:
        x2 = x
    if x2.stride(1) != 1 and x2.shape[1] > 1:
        raise DGLError("gather_rows: x must have unit inner stride")
    n_out = int(idx.shape[0]) if idx is not None else int(x2.shape[0])
    if idx is not None:
        idx = idx.contiguous()
        if idx.dtype != torch.int64:
            raise DGLError("gather_rows: ids must be int64")
    dev = out.device if out is not None else (idx.device if idx is not None else x.device)
    ret = out
    if out is None:
        ret = torch.empty((n_out,) + tail, dtype=x.dtype, device=dev)
    o2 = ret.reshape(n_out, -1) if ret.is_contiguous() else ret
    if o2.dim() != 2 or (o2.stride(1) != 1 and o2.shape[1] > 1):
        raise DGLError("gather_rows: out must be 2-D with unit inner stride")
    es = x2.element_size()
    row_bytes = int(x2.shape[1]) * es
    x_ld = int(x2.stride(0)) * es if x2.shape[0] > 1 else row_bytes
    o_ld = int(o2.stride(0)) * es if n_out > 1 else row_bytes
    bad = _BAD_FLAGS.get(dev)
    if bad is None:
        bad = _BAD_FLAGS[dev] = torch.zeros(1, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        _lib.check_call(lib.dglb200_gather_rows_strided(n_out, _ptr(idx), int(x2.shape[0]), _ptr(x2), max(x_ld, row_bytes),
                                                        row_bytes, _ptr(o2), max(o_ld, row_bytes), _ptr(bad),
                                                        int(max_ctas), _stream()))
    return ret


def gather_rows_convert(x, idx, out, max_ctas=0):
    """out[i] = x[idx[i]] (or x[i] when idx is None) converted float32 -> bfloat16 (round to nearest
    even) or bfloat16 -> float32 (dglb200_gather_rows_convert).  2-D tensors with unit inner stride;
    `out` may alias a peer GPU's mapped memory.  The reduced-precision wire mode of dist.py."""
    lib = _lib.get_lib()
    _need_cuda(x, idx, out)
    if (x.dtype, out.dtype) == (torch.float32, torch.bfloat16):
        to_bf16 = 1
    elif (x.dtype, out.dtype) == (torch.bfloat16, torch.float32):
        to_bf16 = 0
    else:
        raise DGLError("gather_rows_convert: needs float32 -> bfloat16 or bfloat16 -> float32, got %s -> %s"
                       % (x.dtype, out.dtype))
    if x.dim() != 2 or out.dim() != 2 or x.shape[1] != out.shape[1]:
        raise DGLError("gather_rows_convert: x and out must be 2-D with the same width")
    if (x.shape[1] > 1 and x.stride(1) != 1) or (out.shape[1] > 1 and out.stride(1) != 1):
        raise DGLError("gather_rows_convert: unit inner stride required")
    n_out = int(idx.shape[0]) if idx is not None else int(x.shape[0])
    if int(out.shape[0]) != n_out:
        raise DGLError("gather_rows_convert: out has %d rows, expected %d" % (out.shape[0], n_out))
    if idx is not None:
        idx = idx.contiguous()
        if idx.dtype != torch.int64:
            raise DGLError("gather_rows_convert: ids must be int64")
    feat = int(x.shape[1])
    x_ld = int(x.stride(0)) if x.shape[0] > 1 else feat
    o_ld = int(out.stride(0)) if n_out > 1 else feat
    dev = out.device
    bad = _BAD_FLAGS.get(dev)
    if bad is None:
        bad = _BAD_FLAGS[dev] = torch.zeros(1, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        _lib.check_call(lib.dglb200_gather_rows_convert(n_out, _ptr(idx), int(x.shape[0]), to_bf16, _ptr(x),
                                                        max(x_ld, feat), feat, _ptr(out), max(o_ld, feat),
                                                        _ptr(bad), int(max_ctas), _stream()))
    return out


def gather_rows_multi(sources, seg_ptr, idx, out, max_ctas=0):
    """out[i] = sources[q][idx[i]] for seg_ptr[q] <= i < seg_ptr[q+1] (dglb200_gather_rows_multi):
    one launch that pulls from several (peer) tensors at once.  sources: dense 2-D tensors of equal
    width; seg_ptr: python list of len(sources)+1 offsets; idx int64 CUDA; out dense 2-D."""
    lib = _lib.get_lib()
    _need_cuda(idx, out)
    n_src = len(sources)
    es = out.element_size()
    row_bytes = int(out.shape[1]) * es
    base = (ctypes.c_void_p * n_src)(*[s.data_ptr() if s is not None else 0 for s in sources])
    rows = (ctypes.c_int64 * n_src)(*[int(s.shape[0]) if s is not None else 0 for s in sources])
    seg = (ctypes.c_int64 * (n_src + 1))(*[int(v) for v in seg_ptr])
    dev = out.device
    bad = _BAD_FLAGS.get(dev)
    if bad is None:
        bad = _BAD_FLAGS[dev] = torch.zeros(1, dtype=torch.int32, device=dev)
    if not out.is_contiguous() or not idx.is_contiguous():
        raise DGLError("gather_rows_multi: out and idx must be contiguous")
    with torch.cuda.device(dev):
        _lib.check_call(lib.dglb200_gather_rows_multi(n_src, base, rows, seg, _ptr(idx), row_bytes, _ptr(out),
                                                      _ptr(bad), int(max_ctas), _stream()))
    return out


def _addr(t):
    return int(t.data_ptr()) if t is not None else 0


def peer_push(x, idx, counts, dst_addrs, dst_ld_bytes, row_bytes, signal_addrs=None, signal_value=0,
              counter=None, wait_flags=None, wait_value=0, err=None, to_bf16=False, rotate=0, max_ctas=0):
    """dglb200_peer_push: ONE launch stores rows of `x` (2-D, unit inner stride; a column block is fine)
    into every destination's buffer and then signals the destinations.  counts[q] rows go to destination
    q: rows idx[sum(counts[:q]) ...] of x (idx int64 CUDA, or None: the segment's row i is x row i).
    dst_addrs[q]: device address (usually peer memory) of the segment's first output row; signal_addrs[q]:
    address of the flag word that receives signal_value once all stores are visible."""
    lib = _lib.get_lib()
    _need_cuda(x, idx)
    n_dst = len(counts)
    seg = [0]
    for c in counts:
        seg.append(seg[-1] + int(c))
    seg_arr = (ctypes.c_int64 * (n_dst + 1))(*seg)
    dst_arr = (ctypes.c_void_p * n_dst)(*[int(a) if a else None for a in dst_addrs])
    sig_arr = None
    if signal_addrs is not None:
        sig_arr = (ctypes.c_void_p * n_dst)(*[int(a) if a else None for a in signal_addrs])
    if idx is not None:
        idx = idx.contiguous()
        if idx.dtype != torch.int64:
            raise DGLError("peer_push: ids must be int64")
    if x.dim() != 2 or (x.shape[1] > 1 and x.stride(1) != 1):
        raise DGLError("peer_push: x must be 2-D with unit inner stride")
    es = x.element_size()
    x_ld = int(x.stride(0)) * es if x.shape[0] > 1 else max(int(row_bytes), int(x.shape[1]) * es)
    dev = x.device
    bad = _BAD_FLAGS.get(dev)
    if bad is None:
        bad = _BAD_FLAGS[dev] = torch.zeros(1, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        _lib.check_call(lib.dglb200_peer_push(
            n_dst, seg_arr, dst_arr, sig_arr, _ptr(idx), int(x.shape[0]), _ptr(x), x_ld, int(row_bytes),
            int(dst_ld_bytes), 1 if to_bf16 else 0, _ptr(wait_flags),
            int(wait_flags.numel()) if wait_flags is not None else 0, int(wait_value), int(signal_value),
            _ptr(counter), _ptr(err), _ptr(bad), int(rotate), int(max_ctas), _stream()))


def peer_signal(signal_addrs, value, device):
    """dglb200_peer_signal: stream-ordered store of `value` into each listed flag word."""
    lib = _lib.get_lib()
    n = len(signal_addrs)
    arr = (ctypes.c_void_p * n)(*[int(a) if a else None for a in signal_addrs])
    with torch.cuda.device(device):
        _lib.check_call(lib.dglb200_peer_signal(n, arr, int(value), _stream()))


def peer_wait(flags, value, err=None):
    """dglb200_peer_wait: kernels enqueued after this call on the current stream start only once every
    word of the LOCAL int32 tensor `flags` is >= value."""
    lib = _lib.get_lib()
    _need_cuda(flags)
    with torch.cuda.device(flags.device):
        _lib.check_call(lib.dglb200_peer_wait(_ptr(flags), int(flags.numel()), int(value), _ptr(err), _stream()))


def check_bad_ids():
    """Raise DGLError if any gather / push since the last check saw an out-of-range row id (the kernels
    skip such rows and raise a per-device flag; the reference's F.gather_row raises at once,
    backend/pytorch/tensor.py:108-109).  Reads the flags: a device -> host sync per device."""
    for dev, flag in list(_BAD_FLAGS.items()):
        if int(flag.item()) != 0:
            flag.zero_()
            raise DGLError("gather_rows: a row id was out of range on %s (rows were skipped)" % (dev,))
