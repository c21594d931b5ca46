"""Scheduler: lowers a message-passing call to an executor plan (plan.Prog).

Mirrors python/dgl/runtime/scheduler.py (schedule_update_all :158-198, schedule_snr :111-156,
schedule_pull :309-361, schedule_send :28-56, schedule_recv :58-109, _gen_reduce :428-468,
schedule_apply_nodes :200-232, schedule_apply_edges :234-277,
_gen_send_reduce :470-570, _apply_with_accum :398-426) and python/dgl/runtime/spmv.py
(analyze_v2v_spmv :11-55, gen_v2v_spmv_schedule :83-113, gen_e2v_spmv_schedule :115-132) with the
same op names, the same variable numbering and the same output guarantee: reduced features are
batched in unique-ascending order of the destination node ids (scheduler.py:481-482).

What differs is eligibility and what an SPMV executor is:
  * every builtin (message, reduce) pair is v2v: sum AND max, copy_edge, per-head / element-wise
    edge weights (the reference lowers only sum with scalar weights and sends the rest through
    the e2v incidence SpMM or degree bucketing - results must equal those fallbacks');
  * SPMV / SPMV_WITH_DATA launch kernel (1) of libdgl_b200.so on a device CSR (autograd.aggregate)
    instead of th.sparse.mm on a COO;
  * edge-subset adjacencies are built on the device (unique + relabel + CSR build kernels)
    instead of the per-call CPU COO of build_adj_matrix_uv (spmv.py:154-227);
  * UDF reducers still go through degree bucketing, with the bucket arrays from kernel (4b).
"""
import torch

from . import ops
from .autograd import Adjacency, aggregate
from .base import DGLError
from .function import BuiltinFunction, BundledFunction, EdgeFunction
from .init import zero_initializer
from .udf import EdgeBatch, NodeBatch

_MSG_CODE = {"copy_src": ops.MSG_COPY_SRC, "src_mul_edge": ops.MSG_SRC_MUL_EDGE,
             "copy_edge": ops.MSG_COPY_EDGE}
_RED_CODE = {"sum": ops.RED_SUM, "max": ops.RED_MAX}


def _once(make):
    """Lazy SpMat variable: the adjacency is built (or fetched from the graph's cache) when the
    first SPMV executor runs, like the reference's per-context adjacency lambda
    (ir/executor.py:455-456)."""
    box = []

    def get():
        if not box:
            box.append(make())
        return box[0]
    return get


def _is_iterable(x):
    return isinstance(x, (list, tuple))


def _standardize(func, what):
    """scheduler.py:362-396: a UDF stays callable, builtins become a list."""
    if _is_iterable(func):
        for fn in func:
            if not isinstance(fn, BuiltinFunction):
                raise DGLError("If specify multiple message/reduce functions, "
                               "all of them must be builtin")
        return list(func)
    if isinstance(func, BuiltinFunction):
        return [func]
    if not callable(func):
        raise DGLError("User-defined %s function must be callable. Got: %s" % (what, str(func)))
    return func


def _flat2(t):
    return t.reshape(t.shape[0], -1) if t.dim() != 2 else t


def _rows(d, idx):
    """READ_ROW: every column of a feature dict at rows idx (None = all rows)."""
    if idx is None:
        return dict(d)
    return {k: v[idx] for k, v in d.items()}


def _write_rows(d, idx, new, total_rows, inplace):
    """WRITE_ROW_ / WRITE_ROW_INPLACE_ (ir/executor.py:936-1060; Column.update frame.py:118-149):
    out-of-place by default so autograd sees a fresh tensor."""
    for k, v in new.items():
        if idx is None:
            d[k] = v
            continue
        if k not in d:
            # new column: rows that are not written come from the column's initializer, zero unless
            # the user set one (Frame.add_column, frame.py:317-337; init.py:47-61)
            init = d.get_initializer(k) if hasattr(d, "get_initializer") else zero_initializer
            base = init((total_rows,) + tuple(v.shape[1:]), v.dtype, v.device, slice(0, total_rows))
            d[k] = base.index_copy(0, idx, v) if not inplace else base.index_copy_(0, idx, v)
        elif inplace:
            d[k].index_copy_(0, idx, v)
        else:
            d[k] = d[k].index_copy(0, idx, v)


# ---- eligibility (spmv.py:11-81 widened) ------------------------------------------------------

def analyze_v2v(g, mfunc, rfunc):
    """Pair each reducer with the message builtin that produces its field.  Returns
    (pairs, mfunc_left, rfunc_left) like analyze_v2v_spmv; a pair is (mfn, rfn, w_width)."""
    pairs, mleft, rleft = [], [], []
    fld2m = {fn.out_field: fn for fn in mfunc}
    touched = set()
    for rfn in rfunc:
        mfld = rfn.msg_field
        if mfld not in fld2m:
            raise DGLError('Reduce function requires message field "%s",'
                           ' but no message function generates it.' % mfld)
        mfn = fld2m[mfld]
        form = _kernel_form(g, mfn) if rfn.name in _RED_CODE else None
        if form is not None:
            pairs.append((mfn, rfn, form))
        else:
            if mfld not in touched:
                touched.add(mfld)
                mleft.append(mfn)
            rleft.append(rfn)
    return pairs, mleft, rleft


def _kernel_form(g, mfn):
    """w_width (>= 1) when kernel (1) computes this message builtin, else None."""
    if mfn.name == "copy_src":
        return 1 if mfn.src_field in g.ndata else None
    if mfn.name == "copy_edge":
        return 1 if mfn.edge_field in g.edata else None
    if mfn.name == "src_mul_edge":
        if mfn.src_field not in g.ndata or mfn.edge_field not in g.edata:
            return None
        return mfn.kernel_form(g)
    return None


# ---- executors --------------------------------------------------------------------------------

def _spmv_fn(msg_code, red_code, w_width):
    """Body of an SPMV / SPMV_WITH_DATA executor (ir/executor.py:452-473, 535-566): flatten the
    trailing feature dims, run kernel (1), restore the shape."""
    def run_copy_src(adj, x):
        adj = adj() if callable(adj) else adj
        y = aggregate(adj, msg_code, red_code, x=_flat2(x.reshape(x.shape[0], -1)))
        return y.reshape((adj.n_dst,) + tuple(x.shape[1:]))

    def run_copy_edge(adj, w):
        adj = adj() if callable(adj) else adj
        y = aggregate(adj, msg_code, red_code, w=w.reshape(w.shape[0], -1))
        return y.reshape((adj.n_dst,) + tuple(w.shape[1:]))

    def run_mul(adj, w, x):
        adj = adj() if callable(adj) else adj
        rank = max(x.dim(), w.dim())
        out_tail = tuple(x.shape[1:]) + (1,) * (rank - x.dim())
        y = aggregate(adj, msg_code, red_code, x=x.reshape(x.shape[0], -1),
                      w=w.reshape(w.shape[0], -1), w_width=w_width)
        return y.reshape((adj.n_dst,) + out_tail)

    if msg_code == ops.MSG_COPY_SRC:
        return run_copy_src
    if msg_code == ops.MSG_COPY_EDGE:
        return run_copy_edge
    return run_mul


def _read_edge(prog, var_ef, var_eid, field):
    """READ(ef, eid, "field") (ir/executor.py:268-272): eid None = every edge, no gather."""
    return prog.issue("READ", lambda ef, eid, f: ef[f] if eid is None else ef[f][eid],
                      [var_ef, var_eid, prog.var("str", field)], "feat")


def gen_v2v_schedule(prog, adj_var, pairs, var_nf, var_ef, var_eid, var_out):
    """gen_v2v_spmv_schedule (spmv.py:83-113)."""
    for mfn, rfn, w_width in pairs:
        msg, red = _MSG_CODE[mfn.name], _RED_CODE[rfn.name]
        if mfn.name == "src_mul_edge":
            ftedge = _read_edge(prog, var_ef, var_eid, mfn.edge_field)
            ftsrc = prog.issue("READ_COL", lambda nf, f: nf[f], [var_nf, prog.var("str", mfn.src_field)], "feat")
            ftdst = prog.issue("SPMV_WITH_DATA", _spmv_fn(msg, red, w_width), [adj_var, ftedge, ftsrc], "feat")
        elif mfn.name == "copy_edge":
            ftedge = _read_edge(prog, var_ef, var_eid, mfn.edge_field)
            ftdst = prog.issue("SPMV", _spmv_fn(msg, red, 1), [adj_var, ftedge], "feat")
        else:
            ftsrc = prog.issue("READ_COL", lambda nf, f: nf[f], [var_nf, prog.var("str", mfn.src_field)], "feat")
            ftdst = prog.issue("SPMV", _spmv_fn(msg, red, 1), [adj_var, ftsrc], "feat")
        prog.issue("WRITE_COL_", lambda out, f, val: out.__setitem__(f, val),
                   [var_out, prog.var("str", rfn.out_field), ftdst])


def gen_e2v_schedule(prog, adj_var, rfuncs, var_mf, var_out):
    """gen_e2v_spmv_schedule (spmv.py:115-132): builtin reduce over materialised messages.  The
    reference multiplies by an N x E incidence matrix; here the same in-edge CSR gathers message
    rows through its edge/message ids (COPY_EDGE form of kernel 1)."""
    for rfn in rfuncs:
        ftmsg = prog.issue("READ_COL", lambda mf, f: mf[f], [var_mf, prog.var("str", rfn.msg_field)], "feat")
        ftdst = prog.issue("SPMV", _spmv_fn(ops.MSG_COPY_EDGE, _RED_CODE[rfn.name], 1), [adj_var, ftmsg], "feat")
        prog.issue("WRITE_COL_", lambda out, f, val: out.__setitem__(f, val),
                   [var_out, prog.var("str", rfn.out_field), ftdst])


def _gen_send(prog, g, var_nf, var_ef, var_u, var_v, var_eid, mfunc):
    """_gen_send (scheduler.py:572-583): materialise per-edge src / dst / edge data, run the UDF."""
    fdsrc = prog.issue("READ_ROW", _rows, [var_nf, var_u], "dict")
    fddst = prog.issue("READ_ROW", _rows, [var_nf, var_v], "dict")
    fdedge = prog.issue("READ_ROW", _rows, [var_ef, var_eid], "dict")

    def wrapper(src_data, edge_data, dst_data):
        eid = var_eid.data
        if eid is None:
            eid = torch.arange(g.number_of_edges(), device=var_u.data.device)
        return mfunc(EdgeBatch(g, (var_u.data, var_v.data, eid), src_data, edge_data, dst_data))

    return prog.issue("EDGE_UDF", lambda fn, s, e, d: fn(s, e, d),
                      [prog.var("func", wrapper), fdsrc, fdedge, fddst], "dict")


def gen_degree_bucketing_schedule(prog, g, rfunc, vids, recv_nodes, var_nf, var_mf, var_out, mids=None):
    """gen_degree_bucketing_schedule (runtime/degree_bucketing.py:13-84) with the bucket arrays
    from kernel (4b) in canonical order.  Message ids are 0..M-1 for messages materialised by the
    same call (scheduler.py:566) and the edge ids for messages parked by send (scheduler.py:466)."""
    n_msg = int(vids.shape[0])
    if mids is None:
        mids = torch.arange(n_msg, dtype=torch.int64, device=vids.device)
    degs, nids, nsec, bmids, msec = ops.degree_bucketing(mids, vids, recv_nodes, g.number_of_nodes())
    degs = degs.tolist()
    node_bkts = list(torch.split(nids, nsec.tolist()))
    msg_bkts = list(torch.split(bmids, msec.tolist()))
    zero_nodes = None
    if degs and degs[-1] == 0:          # _process_buckets :170-176
        degs = degs[:-1]
        zero_nodes = node_bkts[-1]
        node_bkts = node_bkts[:-1]
    idx_list, fd_list = [], []
    for deg, vbkt, mid in zip(degs, node_bkts, msg_bkts):
        def per_bkt(node_data, mail_data, deg=deg, vbkt=vbkt):
            mail = {k: v.reshape((vbkt.shape[0], deg) + tuple(v.shape[1:])) for k, v in mail_data.items()}
            return rfunc(NodeBatch(g, vbkt, node_data, mail))
        v_vbkt = prog.var("idx", vbkt)
        v_mid = prog.var("idx", mid)
        v_fn = prog.var("func", per_bkt)
        fdvb = prog.issue("READ_ROW", _rows, [var_nf, v_vbkt], "dict")
        fdmail = prog.issue("READ_ROW", _rows, [var_mf, v_mid], "dict")
        fdvb = prog.issue("NODE_UDF", lambda fn, nd, md: fn(nd, md), [v_fn, fdvb, fdmail], "dict")
        idx_list.append(v_vbkt)
        fd_list.append(fdvb)
    if zero_nodes is not None:
        v0 = prog.var("idx", zero_nodes)

        def new_dict(_out, idx, scheme):
            # NewDictExecutor (ir/executor.py:765-778): rows from the node frame's initializers
            # (the reference's tmp frame is frame_like(node_frame), scheduler.py:443)
            return {k: g.ndata.get_initializer(k)((idx.shape[0],) + tuple(v.shape[1:]), v.dtype, v.device, idx)
                    for k, v in scheme.items()}
        zero_feat = prog.issue("NEW_DICT", new_dict, [var_out, v0, fd_list[0]], "dict")
        idx_list.append(v0)
        fd_list.append(zero_feat)
    order = torch.sort(torch.cat([v.data for v in idx_list]))[1]
    var_order = prog.var("idx", order)

    def merge(order, *fds):               # MergeRowExecutor (ir/executor.py:626-634)
        return {k: torch.cat([fd[k] for fd in fds], 0)[order] for k in fds[0]}
    reduced = prog.issue("MERGE_ROW", merge, [var_order] + fd_list, "dict")
    prog.issue("WRITE_DICT_", lambda out, new: out.update(new), [var_out, reduced])


def _gen_send_reduce(prog, g, message_func, reduce_func, var_eid, recv_nodes, n_recv,
                     uv_getter, adj_creator):
    """_gen_send_reduce (scheduler.py:470-570).  Returns the FeatDict var of reduced features,
    rows in unique-ascending destination order."""
    var_nf = prog.var("dict", g.ndata, "nf")
    var_ef = prog.var("dict", g.edata, "ef")
    mfunc = _standardize(message_func, "message")
    rfunc = _standardize(reduce_func, "reduce")
    m_is_list, r_is_list = _is_iterable(mfunc), _is_iterable(rfunc)
    var_out = prog.var("dict", {})
    adj_var = None
    if m_is_list and r_is_list:
        pairs, mfunc, rfunc = analyze_v2v(g, mfunc, rfunc)
        adj_var = prog.var("spmat", _once(adj_creator))
        gen_v2v_schedule(prog, adj_var, pairs, var_nf, var_ef, var_eid, var_out)
        if len(mfunc) == 0:
            return var_out
    if m_is_list:
        mfunc = BundledFunction(mfunc)
    var_u, var_v = uv_getter()
    var_mf = _gen_send(prog, g, var_nf, var_ef, var_u, var_v, var_eid, mfunc)
    if r_is_list:
        kernel_r = [r for r in rfunc if r.name in _RED_CODE]
        rfunc = [r for r in rfunc if r.name not in _RED_CODE]
        if adj_var is None:
            adj_var = prog.var("spmat", _once(adj_creator))
        gen_e2v_schedule(prog, adj_var, kernel_r, var_mf, var_out)
        if len(rfunc) == 0:
            return var_out
        rfunc = BundledFunction(rfunc)
    nodes = recv_nodes
    if nodes is None:
        nodes = torch.arange(n_recv, dtype=torch.int64, device=var_v.data.device)
    gen_degree_bucketing_schedule(prog, g, rfunc, var_v.data, nodes, var_nf, var_mf, var_out)
    return var_out


def _apply_with_accum(prog, g, var_nodes, var_nf, var_accum, apply_func):
    """_apply_with_accum (scheduler.py:398-426)."""
    if not apply_func:
        return var_accum
    v_nf = prog.issue("READ_ROW", _rows, [var_nf, var_nodes], "dict")
    v_nf = prog.issue("UPDATE_DICT", lambda a, b: dict(a, **b), [v_nf, var_accum], "dict")

    def wrapper(node_data):
        return apply_func(NodeBatch(g, var_nodes.data, node_data))
    applied = prog.issue("NODE_UDF", lambda fn, nd: fn(nd), [prog.var("func", wrapper), v_nf], "dict")
    return prog.issue("UPDATE_DICT", lambda a, b: dict(a, **b), [var_accum, applied], "dict")


def subset_adjacency(g, u, v, recv_nodes, inverse):
    """Adjacency of an edge subset: rows = destinations relabelled to their position in
    `recv_nodes` (unique ascending), cols = source ids, slot payload = position of the edge in the
    subset (its message id).  Device-side replacement of build_adj_matrix_uv + build_relabel_map
    (spmv.py:154-227, utils.py:321-361)."""
    return Adjacency(int(recv_nodes.shape[0]), g.number_of_nodes(), coo=(inverse, u, None))


# ---- entry points -------------------------------------------------------------------------------

def schedule_update_all(prog, g, message_func, reduce_func, apply_func):
    """schedule_update_all (scheduler.py:158-198)."""
    if g.number_of_edges() == 0:
        if apply_func is not None:
            schedule_apply_nodes(prog, g, None, apply_func, inplace=False)
        return
    var_nf = prog.var("dict", g.ndata, "nf")
    var_recv = prog.var("idx", None, "recv_nodes")     # slice(0, N): ALL
    var_eid = prog.var("idx", None)                    # slice(0, E): ALL

    def uv_getter():
        src, dst = g.edges()
        return prog.var("idx", src), prog.var("idx", dst)
    reduced = _gen_send_reduce(prog, g, message_func, reduce_func, var_eid, None,
                               g.number_of_nodes(), uv_getter, g.adjacency)
    final = _apply_with_accum(prog, g, var_recv, var_nf, reduced, apply_func)
    prog.issue("WRITE_DICT_", lambda nf, new: nf.update(new), [var_nf, final])


def schedule_snr(prog, g, edge_tuples, message_func, reduce_func, apply_func, inplace):
    """schedule_snr (scheduler.py:111-156)."""
    u, v, eid = edge_tuples
    recv_nodes, inverse = ops.unique_sorted(v, g.number_of_nodes())   # sort_1d(unique(v)) + relabel
    var_nf = prog.var("dict", g.ndata, "nf")
    var_u = prog.var("idx", u)
    var_v = prog.var("idx", v)
    var_eid = prog.var("idx", eid)
    var_recv = prog.var("idx", recv_nodes, "recv_nodes")
    reduced = _gen_send_reduce(prog, g, message_func, reduce_func, var_eid, recv_nodes,
                               int(recv_nodes.shape[0]), lambda: (var_u, var_v),
                               lambda: subset_adjacency(g, u, v, recv_nodes, inverse))
    final = _apply_with_accum(prog, g, var_recv, var_nf, reduced, apply_func)
    n = g.number_of_nodes()
    prog.issue("WRITE_ROW_INPLACE_" if inplace else "WRITE_ROW_",
               lambda nf, idx, new: _write_rows(nf, idx, new, n, inplace), [var_nf, var_recv, final])


def schedule_send(prog, g, u, v, eid, message_func):
    """schedule_send (scheduler.py:28-56): materialise the messages of the given edges and write
    them into the message frame at rows eid (None = every edge).  The reference only accepts a UDF
    here (its TODO at :42); builtins are accepted too and run with their UDF-equivalent __call__."""
    mfunc = _standardize(message_func, "message")
    if _is_iterable(mfunc):
        mfunc = BundledFunction(mfunc)
    var_nf = prog.var("dict", g.ndata, "nf")
    var_ef = prog.var("dict", g.edata, "ef")
    var_mf = prog.var("dict", g._msg, "mf")
    var_u = prog.var("idx", u)
    var_v = prog.var("idx", v)
    var_eid = prog.var("idx", eid)
    msg = _gen_send(prog, g, var_nf, var_ef, var_u, var_v, var_eid, mfunc)
    m = g.number_of_edges()
    prog.issue("WRITE_ROW_", lambda mf, idx, val: _write_rows(mf, idx, val, m, False), [var_mf, var_eid, msg])
    # set message indicator to 1 (scheduler.py:55-56); done at schedule time like the reference
    flags = g.msg_index()
    if eid is None:
        flags.fill_(1)
    else:
        flags[eid] = 1


def _gen_reduce(prog, g, reduce_func, edge_tuples, recv_nodes, inverse):
    """_gen_reduce (scheduler.py:428-468): reduce messages that are already in the message frame.
    Builtin reducers run kernel (1) in its COPY_EDGE form over the sub-CSR of the triggered edges
    (the reference multiplies by the n x M incidence matrix of build_inc_matrix_eid,
    spmv.py:249-314, and only for sum); UDF reducers go through degree bucketing with the edge ids
    as message ids."""
    src, dst, eid = edge_tuples
    rfunc = _standardize(reduce_func, "reduce")
    var_msg = prog.var("dict", g._msg, "msg")
    var_nf = prog.var("dict", g.ndata, "nf")
    var_out = prog.var("dict", {})
    if _is_iterable(rfunc):
        kernel_r = [r for r in rfunc if r.name in _RED_CODE]
        rfunc = [r for r in rfunc if r.name not in _RED_CODE]
        # rows = position of the destination in recv_nodes, slot payload = position in `eid`
        adj_var = prog.var("spmat", _once(lambda: subset_adjacency(g, src, dst, recv_nodes, inverse)))
        for rfn in kernel_r:
            ftmsg = prog.issue("READ_COL", lambda mf, f: mf[f][eid], [var_msg, prog.var("str", rfn.msg_field)], "feat")
            ftdst = prog.issue("SPMV", _spmv_fn(ops.MSG_COPY_EDGE, _RED_CODE[rfn.name], 1), [adj_var, ftmsg], "feat")
            prog.issue("WRITE_COL_", lambda out, f, val: out.__setitem__(f, val),
                       [var_out, prog.var("str", rfn.out_field), ftdst])
        if len(rfunc) == 0:
            return var_out
        rfunc = BundledFunction(rfunc)
    gen_degree_bucketing_schedule(prog, g, rfunc, dst, recv_nodes, var_nf, var_msg, var_out, mids=eid)
    return var_out


def schedule_recv(prog, g, recv_nodes, reduce_func, apply_func, inplace):
    """schedule_recv (scheduler.py:58-109)."""
    src, dst, eid = g.in_edges(recv_nodes)
    flags = g.msg_index()
    if eid.numel() > 0:
        keep = flags[eid].nonzero().reshape(-1)          # only edges that hold a message
        src, dst, eid = src[keep], dst[keep], eid[keep]
    if eid.numel() == 0:
        # downgrade to apply_nodes: all recv nodes are 0-degree, or nothing was sent (:81-86)
        if apply_func is not None:
            schedule_apply_nodes(prog, g, recv_nodes, apply_func, inplace)
        return
    nodes, _ = ops.unique_sorted(recv_nodes, g.number_of_nodes(), want_inverse=False)
    inverse = torch.searchsorted(nodes, dst)
    var_nf = prog.var("dict", g.ndata, "nf")
    var_recv = prog.var("idx", nodes, "recv_nodes")
    reduced = _gen_reduce(prog, g, reduce_func, (src, dst, eid), nodes, inverse)
    final = _apply_with_accum(prog, g, var_recv, var_nf, reduced, apply_func)
    n = g.number_of_nodes()
    prog.issue("WRITE_ROW_INPLACE_" if inplace else "WRITE_ROW_",
               lambda nf, idx, new: _write_rows(nf, idx, new, n, inplace), [var_nf, var_recv, final])
    # the messages are consumed (:106-109)
    flags[eid] = 0
    if not bool(flags.any()):
        prog.issue("CLEAR_FRAME_", lambda mf: mf.clear(), [prog.var("dict", g._msg, "mf")])


def schedule_pull(prog, g, pull_nodes, message_func, reduce_func, apply_func, inplace):
    """schedule_pull (scheduler.py:309-361): zero-in-degree pull nodes get all-zero rows from the
    SpMM (rows without slots) and initializer rows on the bucketing path."""
    u, v, eid = g.in_edges(pull_nodes)
    if eid.numel() == 0:
        if apply_func is not None:
            schedule_apply_nodes(prog, g, pull_nodes, apply_func, inplace)
        return
    nodes, _ = ops.unique_sorted(pull_nodes, g.number_of_nodes(), want_inverse=False)
    # relabel every edge destination to its position in `nodes` (nodes is sorted)
    inverse = torch.searchsorted(nodes, v)
    var_nf = prog.var("dict", g.ndata, "nf")
    var_pull = prog.var("idx", nodes, "pull_nodes")
    var_u = prog.var("idx", u)
    var_v = prog.var("idx", v)
    var_eid = prog.var("idx", eid)
    reduced = _gen_send_reduce(prog, g, message_func, reduce_func, var_eid, nodes,
                               int(nodes.shape[0]), lambda: (var_u, var_v),
                               lambda: subset_adjacency(g, u, v, nodes, inverse))
    final = _apply_with_accum(prog, g, var_pull, var_nf, reduced, apply_func)
    n = g.number_of_nodes()
    prog.issue("WRITE_ROW_INPLACE_" if inplace else "WRITE_ROW_",
               lambda nf, idx, new: _write_rows(nf, idx, new, n, inplace), [var_nf, var_pull, final])


def schedule_apply_nodes(prog, g, v, apply_func, inplace):
    """schedule_apply_nodes (scheduler.py:200-232); v None = all nodes."""
    var_nf = prog.var("dict", g.ndata, "nf")
    var_v = prog.var("idx", v)
    v_nf = prog.issue("READ_ROW", _rows, [var_nf, var_v], "dict")

    def wrapper(node_data):
        return apply_func(NodeBatch(g, v, node_data))
    applied = prog.issue("NODE_UDF", lambda fn, nd: fn(nd), [prog.var("func", wrapper), v_nf], "dict")
    n = g.number_of_nodes()
    prog.issue("WRITE_ROW_INPLACE_" if inplace else "WRITE_ROW_",
               lambda nf, idx, new: _write_rows(nf, idx, new, n, inplace), [var_nf, var_v, applied])


def schedule_apply_edges(prog, g, u, v, eid, apply_func, inplace):
    """schedule_apply_edges (scheduler.py:234-277); eid None = all edges.  A builtin edge function
    over all edges is ONE SDDMM launch on the cached in-edge CSR (no E-row gathers, no UDF)."""
    var_nf = prog.var("dict", g.ndata, "nf")
    var_ef = prog.var("dict", g.edata, "ef")
    if isinstance(apply_func, EdgeFunction) and eid is None:
        from . import edge_ops
        fn = apply_func
        adj_var = prog.var("spmat", _once(g.adjacency))
        read = lambda f: (prog.issue("READ_COL", lambda nf, k: nf[k], [var_nf, prog.var("str", f)], "feat")  # noqa: E731
                          if f is not None else prog.var("feat", None))
        lhs, rhs = read(fn.lhs_field), read(fn.rhs_field)
        out = prog.issue("SDDMM", lambda adj, a, b: edge_ops.sddmm(adj(), fn.name, a, b), [adj_var, lhs, rhs], "feat")
        prog.issue("WRITE_COL_", lambda ef, f, val: ef.__setitem__(f, val),
                   [var_ef, prog.var("str", fn.out_field), out])
        return
    var_u = prog.var("idx", u)
    var_v = prog.var("idx", v)
    var_eid = prog.var("idx", eid)
    fdsrc = prog.issue("READ_ROW", _rows, [var_nf, var_u], "dict")
    fddst = prog.issue("READ_ROW", _rows, [var_nf, var_v], "dict")
    fdedge = prog.issue("READ_ROW", _rows, [var_ef, var_eid], "dict")

    def wrapper(src_data, edge_data, dst_data):
        e = eid if eid is not None else torch.arange(g.number_of_edges(), device=u.device)
        return apply_func(EdgeBatch(g, (u, v, e), src_data, edge_data, dst_data))
    new = prog.issue("EDGE_UDF", lambda fn, s, e, d: fn(s, e, d),
                     [prog.var("func", wrapper), fdsrc, fdedge, fddst], "dict")
    m = g.number_of_edges()
    prog.issue("WRITE_ROW_INPLACE_" if inplace else "WRITE_ROW_",
               lambda ef, idx, val: _write_rows(ef, idx, val, m, inplace), [var_ef, var_eid, new])
