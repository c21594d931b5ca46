"""DGLGraph: host-side mirror of the reference's message-passing entry points.

Same method names, argument meaning and ordering guarantees as python/dgl/graph.py
(update_all :2398, send_and_recv :2089, pull :2198, push :2301, apply_nodes :1800,
apply_edges :1854) for the builtin-function path; graph mutation / queries beyond what that path
needs are out of scope.  Structure lives on the device as an edge list in edge-id
order plus cached in-/out-edge CSRs (built by kernel 4); features are torch tensors.

Scheduling follows python/dgl/runtime/scheduler.py (see scheduler.py in this package); every
call records its executor plan in `last_prog` (plan.Prog) before running it.
"""
import torch

from . import ops
from .autograd import Adjacency
from .base import ALL, DGLError, is_all
from .init import zero_initializer


class FeatureDict(dict):
    """ndata / edata: field -> tensor with a fixed number of rows (cf. Frame, frame.py:179)."""

    def __init__(self, num_rows_fn, kind):
        super(FeatureDict, self).__init__()
        self._num_rows_fn = num_rows_fn
        self._kind = kind
        self._initializers = {}
        self._default_initializer = zero_initializer

    def set_initializer(self, initializer, column=None):
        """Frame.set_initializer (python/dgl/frame.py:240-256)."""
        if column is None:
            self._default_initializer = initializer
        else:
            self._initializers[column] = initializer

    def get_initializer(self, column=None):
        """Frame.get_initializer (python/dgl/frame.py:225-238); the default is the zero
        initializer the reference falls back to (frame.py:219-223)."""
        return self._initializers.get(column, self._default_initializer)

    def num_rows(self):
        return self._num_rows_fn()

    def __setitem__(self, key, val):
        if not torch.is_tensor(val):
            raise DGLError("%s feature '%s' must be a tensor" % (self._kind, key))
        if val.shape[0] != self._num_rows_fn():
            raise DGLError("Expect number of features to match number of %ss. Got %d and %d instead."
                           % (self._kind, val.shape[0], self._num_rows_fn()))
        super(FeatureDict, self).__setitem__(key, val)


def _as_ids(v, device=None):
    """Normalise an id container to an int64 tensor (utils.Index, python/dgl/utils.py:12-219)."""
    if torch.is_tensor(v):
        t = v.to(torch.int64)
    elif isinstance(v, int):
        t = torch.tensor([v], dtype=torch.int64)
    else:
        t = torch.as_tensor(list(v), dtype=torch.int64)
    if device is not None and t.device != device:
        t = t.to(device)
    return t.reshape(-1)


class DGLGraph(object):
    def __init__(self, graph_data=None, multigraph=True, readonly=False):
        """graph_data: None, a scipy sparse matrix (row = source, column = destination, edge ids in the
        order of its COO form: GraphIndex.from_scipy_sparse_matrix, python/dgl/graph_index.py:752-768) or an
        edge list [(u, v), ...] (from_edge_list, :771-790).  readonly=True freezes the structure after
        construction (the reference's ImmutableGraph; needed by the neighbour sampler)."""
        self._n = 0
        self._src = torch.empty(0, dtype=torch.int64)
        self._dst = torch.empty(0, dtype=torch.int64)
        self._e = 0
        self._device = None
        self._adj = None            # Adjacency of the whole graph (cache)
        self._readonly = readonly
        self.ndata = FeatureDict(self.number_of_nodes, "node")
        self.edata = FeatureDict(self.number_of_edges, "edge")
        # message frame + per-edge "has a pending message" flag (graph.py:191-196 in the reference)
        self._msg = FeatureDict(self.number_of_edges, "message")
        self._msg_index = None
        self._message_func = None
        self._reduce_func = None
        self._apply_node_func = None
        self._apply_edge_func = None
        self.last_prog = None
        self.execute = True         # False: only record the plan in last_prog (tests diff it)
        self._imm = {}              # neighbour type -> (row, col)-sorted CSR (ImmutableGraph::CSR::FromEdges)
        self._frozen = False
        if graph_data is None and readonly:
            raise DGLError("can't create an empty immutable graph")        # graph_index.py:971-972
        if graph_data is not None:
            self._from_data(graph_data)
        self._frozen = bool(readonly)

    def _from_data(self, data):
        if hasattr(data, "tocoo"):
            coo = data.tocoo()
            n = max(coo.shape[0], coo.shape[1])
            u = torch.as_tensor(coo.row, dtype=torch.int64)
            v = torch.as_tensor(coo.col, dtype=torch.int64)
        else:
            try:
                pairs = [(int(a), int(b)) for a, b in data]
            except Exception:
                raise DGLError('Error while creating graph from input of type "%s".' % type(data))
            u = torch.tensor([p[0] for p in pairs], dtype=torch.int64)
            v = torch.tensor([p[1] for p in pairs], dtype=torch.int64)
            n = int(max(int(u.max()), int(v.max())) + 1) if pairs else 0
        self.add_nodes(n)
        if u.numel():
            self.add_edges(u, v)

    def is_readonly(self):
        return self._readonly

    def immutable_csr(self, neighbor_type="in"):
        """The (row, column)-sorted CSR of the reference's ImmutableGraph (CSR::FromEdges,
        src/graph/immutable_graph.cc:206-237; "in": keyed by destination, else by source), cached."""
        key = "in" if neighbor_type == "in" else "out"
        if key not in self._imm:
            self._need_device()
            src, dst = self.edges()
            r, c = (dst, src) if key == "in" else (src, dst)
            self._imm[key] = ops.csr_build(r, c, self._n, self._n, order=1)
        return self._imm[key]

    # ---- structure -----------------------------------------------------------------------
    def add_nodes(self, num):
        if self._frozen:
            raise DGLError("Readonly graph. Mutation is not allowed.")
        self._n += int(num)
        self.clear_cache()

    def add_edge(self, u, v):
        self.add_edges([u], [v])

    def add_edges(self, u, v):
        """Append edges; edge ids are assigned in insertion order (Graph::AddEdges, graph.cc:66-97)."""
        if self._frozen:
            raise DGLError("Readonly graph. Mutation is not allowed.")
        dev = self._device if self._device is not None else (u.device if torch.is_tensor(u) else None)
        u = _as_ids(u, dev)
        v = _as_ids(v, u.device)
        if u.numel() == 1 and v.numel() > 1:
            u = u.expand(v.numel())
        if v.numel() == 1 and u.numel() > 1:
            v = v.expand(u.numel())
        if u.numel() != v.numel():
            raise DGLError("add_edges: u and v have different lengths")
        if self._e > 0 and self._src is None:
            self._src, self._dst = self.edges()     # the edge list was dropped: recover it from the CSRs
        if self._e == 0:
            self._src, self._dst = u.contiguous(), v.contiguous()
        else:
            self._src = torch.cat([self._src.to(u.device), u])
            self._dst = torch.cat([self._dst.to(u.device), v])
        self._e = int(self._src.shape[0])
        if u.is_cuda:
            self._device = u.device
        self.clear_cache()
        self._reset_messages()

    def to(self, device):
        """Move the structure to `device` (features are moved by the user, as in the reference)."""
        device = torch.device(device)
        if self._e > 0 and self._src is None:
            self._src, self._dst = self.edges()     # dropped edge list: recover before the cache goes
        self._device = device
        if self._src is not None:
            self._src = self._src.to(device)
            self._dst = self._dst.to(device)
        self.clear_cache()
        self._reset_messages()
        return self

    def number_of_nodes(self):
        return self._n

    def number_of_edges(self):
        return self._e

    def __len__(self):
        return self._n

    def clear_cache(self):
        """Drop the cached device CSRs (GraphIndex.clear_cache, graph_index.py:113-115)."""
        self._adj = None
        self._imm = {}

    def _reset_messages(self):
        """Structure changed: pending messages are dropped (the reference appends zero rows to
        _msg_index / _msg_frame on add_edges, graph.py:308-309,368-369; here the flags restart)."""
        self._msg.clear()
        self._msg_index = None

    def msg_index(self):
        """int8[E] device flags: 1 = the edge holds a message that no recv has consumed."""
        if self._msg_index is None or self._msg_index.shape[0] != self._e:
            self._msg_index = torch.zeros(self._e, dtype=torch.int8, device=self._need_device())
        return self._msg_index

    def set_n_initializer(self, initializer, field=None):
        """python/dgl/graph.py:1590-1618."""
        self.ndata.set_initializer(initializer, field)

    def set_e_initializer(self, initializer, field=None):
        """python/dgl/graph.py:1620-1648."""
        self.edata.set_initializer(initializer, field)

    def _need_device(self):
        if self._device is None or self._device.type != "cuda":
            raise DGLError("the graph structure is not on a CUDA device: call g.to('cuda') first "
                           "(dgl_b200 has no CPU path)")
        return self._device

    def adjacency(self):
        """Whole-graph Adjacency (cached): rows = dst, cols = src, edge-id order inside a row, i.e.
        Graph::GetAdj(false, "csr") (src/graph/graph.cc:525-549)."""
        if self._adj is None:
            self._need_device()
            if self._src is None:
                raise DGLError("edge list was dropped and the CSR cache was cleared")
            self._adj = Adjacency(self._n, self._n, coo=(self._dst, self._src, None))
        return self._adj

    def in_csr(self):
        return self.adjacency().csr_in

    def out_csr(self):
        return self.adjacency().csr_out

    def set_csr(self, csr_in, csr_out):
        """Install prebuilt CSRs (benchmarks that build / slice the structure themselves)."""
        self._adj = Adjacency(csr_in.n_rows, csr_in.n_cols, csr_in=csr_in, csr_out=csr_out)

    def drop_coo(self):
        """Free the int64 edge list once both CSRs exist (16 B/edge); subset ops that need
        (u, v) pairs recover them from the CSRs."""
        self.adjacency().drop_coo()
        self._src = self._dst = None

    def adjacency_csr(self, transpose=False):
        """(indptr, indices, eid) int64, as _CAPI_DGLGraphGetAdj(handle, transpose, "csr") returns
        (src/graph/graph_apis.cc:474-482)."""
        csr = self.out_csr() if transpose else self.in_csr()
        return csr.indptr, ops.widen(csr.indices), ops.widen(csr.eids)

    def edges(self):
        """(src, dst) in edge-id order (Graph::Edges, graph.cc:368-409, sorted=False)."""
        if self._src is None:
            csr = self.in_csr()
            src = torch.empty(self._e, dtype=torch.int64, device=self._device)
            dst = torch.empty(self._e, dtype=torch.int64, device=self._device)
            e = csr.eids.long()
            src[e] = csr.indices.long()
            dst[e] = ops.csr_row_ids(csr)
            return src, dst
        return self._src, self._dst

    def in_degrees(self, v=ALL):
        """Graph::InDegrees (graph.cc:412-424)."""
        if is_all(v):
            return ops.degrees(self.in_csr())
        return ops.degrees_of(self.in_csr(), _as_ids(v, self._need_device()))

    def out_degrees(self, v=ALL):
        if is_all(v):
            return ops.degrees(self.out_csr())
        return ops.degrees_of(self.out_csr(), _as_ids(v, self._need_device()))

    # ---- edge queries used by the subset entry points ---------------------------------------
    @staticmethod
    def _rows_of(csr, nodes):
        """Slots of the CSR rows `nodes` (in the order given): (row position, slot) int64 - one scan +
        one expand kernel on the device (dglb200_csr_select_*)."""
        return ops.csr_select(csr, nodes, want=("which", "slot"))

    def in_edges(self, v):
        """(u, v, eid) of the in-edges of nodes v, per node in edge-id order
        (Graph::InEdges(IdArray), graph.cc:294-319)."""
        v = _as_ids(v, self._need_device())
        csr = self.in_csr()
        which, slot = self._rows_of(csr, v)
        return csr.indices[slot].long(), v[which], csr.eids[slot].long()

    def out_edges(self, u):
        """(u, v, eid) of the out-edges of nodes u (Graph::OutEdges(IdArray), graph.cc:340-365)."""
        u = _as_ids(u, self._need_device())
        csr = self.out_csr()
        which, slot = self._rows_of(csr, u)
        return u[which], csr.indices[slot].long(), csr.eids[slot].long()

    def find_edges(self, eid):
        """(u, v, eid) of the given edge ids (Graph::FindEdges, graph.cc:252-269)."""
        eid = _as_ids(eid, self._need_device())
        src, dst = self.edges()
        return src[eid], dst[eid], eid

    def edge_ids(self, u, v):
        """All edges between the pairs (u[i], v[i]) - (src, dst, eid), grouped by pair, ascending
        edge id inside a pair (Graph::EdgeIds, graph.cc:205-249)."""
        dev = self._need_device()
        u = _as_ids(u, dev)
        v = _as_ids(v, dev)
        if u.numel() == 1 and v.numel() > 1:
            u = u.expand(v.numel())
        if v.numel() == 1 and u.numel() > 1:
            v = v.expand(u.numel())
        csr = self.out_csr()                       # row u: (dst, eid) in ascending edge id
        which, slot = self._rows_of(csr, u)
        hit = csr.indices[slot].long() == v[which]
        which, slot = which[hit], slot[hit]
        return u[which], v[which], csr.eids[slot].long()

    # ---- default functions (graph.py:1700-1798 register_*_func) ------------------------------
    def register_message_func(self, func):
        self._message_func = func

    def register_reduce_func(self, func):
        self._reduce_func = func

    def register_apply_node_func(self, func):
        self._apply_node_func = func

    def register_apply_edge_func(self, func):
        self._apply_edge_func = func

    def _defaults(self, message_func, reduce_func, apply_node_func):
        if message_func == "default":
            message_func = self._message_func
        if reduce_func == "default":
            reduce_func = self._reduce_func
        if apply_node_func == "default":
            apply_node_func = self._apply_node_func
        assert message_func is not None
        assert reduce_func is not None
        return message_func, reduce_func, apply_node_func

    # ---- message passing entry points --------------------------------------------------------
    def _run(self, build):
        from .plan import Prog
        prog = Prog()
        build(prog)
        self.last_prog = prog
        if self.execute:
            prog.run()

    def update_all(self, message_func="default", reduce_func="default", apply_node_func="default"):
        """python/dgl/graph.py:2398-2441."""
        from . import scheduler
        m, r, a = self._defaults(message_func, reduce_func, apply_node_func)
        self._run(lambda prog: scheduler.schedule_update_all(prog, self, m, r, a))

    def _resolve_edges(self, edges):
        """(u, v, eid) of an edge specification: ALL, a (u, v) pair (every parallel edge between
        the pairs, Graph::EdgeIds) or an edge-id container (graph.py:1962-1974)."""
        if is_all(edges):
            u, v = self.edges()
            return u, v, None
        if isinstance(edges, tuple):
            return self.edge_ids(edges[0], edges[1])
        return self.find_edges(edges)

    def send(self, edges=ALL, message_func="default"):
        """python/dgl/graph.py:1926-1985: compute messages on `edges` and park them in the message
        frame until a recv consumes them; a later send on the same edge overwrites."""
        from . import scheduler
        if message_func == "default":
            message_func = self._message_func
        assert message_func is not None
        self._need_device()
        u, v, eid = self._resolve_edges(edges)
        if eid is not None and eid.numel() == 0:
            return   # no edge to be triggered
        if self._e == 0:
            return
        self._run(lambda prog: scheduler.schedule_send(prog, self, u, v, eid, message_func))

    def recv(self, v=ALL, reduce_func="default", apply_node_func="default", inplace=False):
        """python/dgl/graph.py:1987-2087: reduce the pending messages on the in-edges of `v`
        (consuming them), then apply."""
        from . import scheduler
        if reduce_func == "default":
            reduce_func = self._reduce_func
        if apply_node_func == "default":
            apply_node_func = self._apply_node_func
        assert reduce_func is not None
        dev = self._need_device()
        if is_all(v):
            v = torch.arange(self._n, dtype=torch.int64, device=dev)
        v = _as_ids(v, dev)
        if v.numel() == 0:
            return   # no vertex to be triggered
        self._run(lambda prog: scheduler.schedule_recv(prog, self, v, reduce_func, apply_node_func,
                                                       inplace))

    def send_and_recv(self, edges, message_func="default", reduce_func="default",
                      apply_node_func="default", inplace=False):
        """python/dgl/graph.py:2089-2196.  `edges` is an edge-id container or a (u, v) pair."""
        from . import scheduler
        m, r, a = self._defaults(message_func, reduce_func, apply_node_func)
        if isinstance(edges, tuple):
            u, v, eid = self.edge_ids(edges[0], edges[1])
        else:
            u, v, eid = self.find_edges(edges)
        if eid.numel() == 0:
            return   # no edge to be triggered (graph.py:2179-2181)
        self._run(lambda prog: scheduler.schedule_snr(prog, self, (u, v, eid), m, r, a, inplace))

    def pull(self, v, message_func="default", reduce_func="default", apply_node_func="default",
             inplace=False):
        """python/dgl/graph.py:2198-2299."""
        from . import scheduler
        m, r, a = self._defaults(message_func, reduce_func, apply_node_func)
        v = _as_ids(v, self._need_device())
        if v.numel() == 0:
            return
        self._run(lambda prog: scheduler.schedule_pull(prog, self, v, m, r, a, inplace))

    def push(self, u, message_func="default", reduce_func="default", apply_node_func="default",
             inplace=False):
        """python/dgl/graph.py:2301-2396: send along the out-edges of u, then receive."""
        from . import scheduler
        m, r, a = self._defaults(message_func, reduce_func, apply_node_func)
        u = _as_ids(u, self._need_device())
        if u.numel() == 0:
            return
        uu, vv, eid = self.out_edges(u)
        if eid.numel() == 0:
            return   # scheduler.py:302-305
        self._run(lambda prog: scheduler.schedule_snr(prog, self, (uu, vv, eid), m, r, a, inplace))

    def prop_nodes(self, nodes_generator, message_func="default", reduce_func="default",
                   apply_node_func="default"):
        """python/dgl/graph.py:2443-2517: pull on every node frontier, in the generator's order."""
        for node_frontier in nodes_generator:
            self.pull(node_frontier, message_func, reduce_func, apply_node_func)

    def prop_edges(self, edges_generator, message_func="default", reduce_func="default",
                   apply_node_func="default"):
        """python/dgl/graph.py:2519-2598: send_and_recv on every edge frontier."""
        for edge_frontier in edges_generator:
            self.send_and_recv(edge_frontier, message_func, reduce_func, apply_node_func)

    def adjacency_coo(self, transpose=False):
        """(idx int64[2E], eid int64[E]) as _CAPI_DGLGraphGetAdj(handle, transpose, "coo") returns:
        idx = [dst | src] in edge-id order ([src | dst] when transposed) and the identity shuffle
        (Graph::GetAdj, src/graph/graph.cc:509-524)."""
        src, dst = self.edges()
        idx = torch.cat([src, dst]) if transpose else torch.cat([dst, src])
        return idx, torch.arange(self._e, dtype=torch.int64, device=idx.device)

    def adjacency_matrix(self, transpose=False, ctx=None):
        """python/dgl/graph.py:2721-2742 -> GraphIndex.adjacency_matrix (graph_index.py:537-585): the
        n x n torch sparse COO the reference hands to users, rows = destinations (sources when
        transposed), one entry of 1.0 per edge in edge-id order (uncoalesced, as the reference's)."""
        if not isinstance(transpose, bool):
            raise DGLError('Expect bool value for "transpose" arg, but got %s.' % (type(transpose)))
        idx, _ = self.adjacency_coo(transpose)
        dat = torch.ones(self._e, dtype=torch.float32, device=idx.device)
        adj = torch.sparse_coo_tensor(idx.view(2, self._e), dat, (self._n, self._n))
        return adj if ctx is None else adj.to(ctx)

    def incidence_matrix(self, typestr, ctx=None):
        """python/dgl/graph.py:2744-2788 -> GraphIndex.incidence_matrix (graph_index.py:587-662):
        n x E sparse COO; 'in': I[v, e] = 1 if v is the destination of e; 'out': if v is its source;
        'both': +1 / -1 for in / out, self loops dropped."""
        src, dst = self.edges()
        eid = torch.arange(self._e, dtype=torch.int64, device=src.device)
        if typestr == "in":
            idx, dat = torch.stack([dst, eid]), torch.ones(self._e, device=src.device)
        elif typestr == "out":
            idx, dat = torch.stack([src, eid]), torch.ones(self._e, device=src.device)
        elif typestr == "both":
            keep = src != dst
            s, d, e = src[keep], dst[keep], eid[keep]
            idx = torch.stack([torch.cat([s, d]), torch.cat([e, e])])
            dat = torch.cat([-torch.ones(s.numel(), device=src.device), torch.ones(s.numel(), device=src.device)])
        else:
            raise DGLError("Invalid incidence matrix type: %s" % str(typestr))
        inc = torch.sparse_coo_tensor(idx, dat, (self._n, self._e))
        return inc if ctx is None else inc.to(ctx)

    def apply_nodes(self, func="default", v=ALL, inplace=False):
        """python/dgl/graph.py:1800-1852."""
        from . import scheduler
        if func == "default":
            func = self._apply_node_func
        v = None if is_all(v) else _as_ids(v, self._device)
        self._run(lambda prog: scheduler.schedule_apply_nodes(prog, self, v, func, inplace))

    def apply_edges(self, func="default", edges=ALL, inplace=False):
        """python/dgl/graph.py:1854-1924."""
        from . import scheduler
        if func == "default":
            func = self._apply_edge_func
        assert func is not None
        from .function import EdgeFunction
        if is_all(edges) and isinstance(func, EdgeFunction):
            u = v = eid = None          # one SDDMM launch over the cached CSR: no edge list needed
        elif is_all(edges):
            u, v = self.edges()
            eid = None
        elif isinstance(edges, tuple):
            u, v, eid = self.edge_ids(edges[0], edges[1])
        else:
            u, v, eid = self.find_edges(edges)
        self._run(lambda prog: scheduler.schedule_apply_edges(prog, self, u, v, eid, func, inplace))
