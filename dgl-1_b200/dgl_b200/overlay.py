"""Drop the engine into an IMPORTED reference (GaiYu0/dgl-1, DGL 0.1.3): the four call sites of the
builtin message-passing path (SURVEY.md 8b "who calls it") are patched in place, user code
(`dgl.function`, `DGLGraph.update_all / send_and_recv / pull / push / recv`) is untouched.

  1. F.sparse_matrix / F.sparse_matrix_indices / F.copy_to / F.spmm
     (python/dgl/backend/pytorch/tensor.py:34-54, 80-87, 138-146): a sparse matrix that lives on an engine
     device is an `EngineAdj` (a CSR pair built by kernel (4)), `F.spmm` on it is kernel (1) with kernel
     (2) / (3) as its autograd - instead of th.sparse.mm on an uncoalesced COO;
  2. SPMVWithDataExecutor.run (python/dgl/runtime/ir/executor.py:535-566): the per-call rebuild of the
     sparse tensor (`F.sparse_matrix(A_data, spidx, shape)`, :545-548) becomes a view that carries the
     edge weights into the src_mul_edge kernel;
  3. the adjacency builders (python/dgl/runtime/spmv.py:134-353 -> GraphIndex.adjacency_matrix /
     incidence_matrix, python/dgl/graph_index.py:537-662) need no patch of their own: they reach the engine
     through 1. (F.copy_to of the index to the device + F.sparse_matrix);
  4. the PackedFunc globals (`runtime.degree_bucketing._CAPI_DGLDegreeBucketing`,
     `graph_index._CAPI_DGLGraph{GetAdj,InEdges_2,OutEdges_2,Edges,InDegrees,OutDegrees,FindEdges,EdgeIds}`) are overridden
     and `graph_index._CAPI_DGLGraphUniformSampling{,2,..,128}` (the neighbour sampler) are overridden
     inside the loaded libdgl.so by dglb200_register_packed_funcs (csrc/packed.cu, packed_graph.cu,
     packed_sampling.cuh).  The reference's Python holds COPIES of the PackedFuncs taken at import
     (DGLFuncGetGlobal, src/runtime/registry.cc:137-147), so install() re-fetches the overridden names into
     the module attributes (and the sampler's dispatch table, graph_index.py:1005-1014).

    import dgl                      # the reference
    from dgl_b200 import overlay
    ov = overlay.install(dgl)       # ... ov.uninstall() restores everything patched in Python

Tensors on a CPU context keep the reference's own path.
"""
import ctypes

import torch

from . import _lib, autograd, ops

# which tensors / contexts belong to the engine (tests point this at "everything" to run the stack on the
# emulated kernels in a container without a GPU)
ON_ENGINE = lambda t: bool(getattr(t, "is_cuda", False))              # noqa: E731
ON_ENGINE_CTX = lambda ctx: getattr(ctx, "type", None) == "cuda"      # noqa: E731


class EngineAdj(object):
    """What F.sparse_matrix returns on an engine device: the CSR pair of `autograd.Adjacency` (+ optional
    per-edge weights in the order of the COO it was made from).  Quacks like the sparse tensor the
    reference's executors handle: `.shape`, and F.copy_to's `.cpu()` / `.cuda()`."""

    def __init__(self, adj, shape, index, weights=None):
        self.adj = adj
        self.shape = tuple(int(s) for s in shape)
        self.index = index            # the (2, nnz) COO index it was built from, in edge order
        self.weights = weights

    def with_weights(self, w):
        return EngineAdj(self.adj, self.shape, self.index, w)

    def cuda(self, *a, **k):
        return self

    def to_torch_sparse(self):
        dat = self.weights if self.weights is not None else torch.ones(self.index.shape[1], device=self.index.device)
        return torch.sparse_coo_tensor(self.index, dat, self.shape)

    def cpu(self):
        return self.to_torch_sparse().cpu()


class Overlay(object):
    def __init__(self, dgl_mod):
        self.dgl = dgl_mod
        self.saved = []
        self.registered = 0

    def _patch(self, obj, name, new):
        self.saved.append((obj, name, getattr(obj, name)))
        setattr(obj, name, new)

    def uninstall(self):
        """Restore the Python-side patches (PackedFunc overrides stay: libdgl has no un-register)."""
        while self.saved:
            obj, name, old = self.saved.pop()
            setattr(obj, name, old)


def _flat_spmm(fn_, B):
    """The reshape rules of SPMVExecutor.run (ir/executor.py:457-472) around a 2-D SpMM."""
    if B.dim() == 1:
        return fn_(B.unsqueeze(1)).squeeze(1)
    if B.dim() > 2:
        feat_shape = tuple(B.shape[1:])
        C = fn_(B.reshape(B.shape[0], -1))
        return C.reshape((C.shape[0],) + feat_shape)
    return fn_(B)


def install(dgl_mod=None, register_packed=True, override_reference=True):
    """Patch an imported reference package.  Returns the Overlay handle."""
    if dgl_mod is None:
        import dgl as dgl_mod          # the reference must be importable (PYTHONPATH, DGL_LIBRARY_PATH)
    F = dgl_mod.backend
    ov = Overlay(dgl_mod)
    orig_sparse_matrix, orig_indices = F.sparse_matrix, F.sparse_matrix_indices
    orig_spmm, orig_copy_to = F.spmm, F.copy_to

    def sparse_matrix(data, index, shape, force_format=False):
        fmt = index[0]
        if fmt == "b200":                                   # SPMVWithDataExecutor.run: same structure, new data
            return index[1].with_weights(data), None
        idx = index[1]
        if fmt == "coo" and ON_ENGINE(idx):
            # rows = idx[0], cols = idx[1], in the order given (edge-id order for whole-graph matrices):
            # slots carry their position as edge id, so no shuffle index is needed
            adj = autograd.Adjacency(shape[0], shape[1], coo=(idx[0].contiguous(), idx[1].contiguous(), None))
            return EngineAdj(adj, shape, idx), None
        return orig_sparse_matrix(data, index, shape, force_format)

    def sparse_matrix_indices(spmat):
        if isinstance(spmat, EngineAdj):
            return ("b200", spmat)
        return orig_indices(spmat)

    def copy_to(inp, ctx):
        if isinstance(inp, EngineAdj):
            return inp if ON_ENGINE_CTX(ctx) else inp.cpu()
        if torch.is_tensor(inp) and inp.is_sparse and ON_ENGINE_CTX(ctx):
            # build_adj_matrix_uv / build_inc_matrix_* make the COO on the CPU and move it per call
            # (runtime/spmv.py:195-227): the index goes to the device, the structure is built there
            idx = orig_copy_to(inp._indices(), ctx)
            return sparse_matrix(None, ("coo", idx), tuple(inp.shape))[0]
        return orig_copy_to(inp, ctx)

    def spmm(x, y):
        if isinstance(x, EngineAdj):
            if x.weights is None:
                return autograd.aggregate(x.adj, ops.MSG_COPY_SRC, ops.RED_SUM, x=y.contiguous())
            w = x.weights.reshape(-1, 1).to(y.dtype)
            return autograd.aggregate(x.adj, ops.MSG_SRC_MUL_EDGE, ops.RED_SUM, x=y.contiguous(), w=w, w_width=1)
        return orig_spmm(x, y)

    ov._patch(F, "sparse_matrix", sparse_matrix)
    ov._patch(F, "sparse_matrix_indices", sparse_matrix_indices)
    ov._patch(F, "copy_to", copy_to)
    ov._patch(F, "spmm", spmm)

    executor = dgl_mod.runtime.ir.executor
    orig_run = executor.SPMVWithDataExecutor.run

    def run_with_data(self):
        B = self.B.data
        spA = self.spA.data(F.context(B))
        if not isinstance(spA, EngineAdj):
            return orig_run(self)
        A_data = self.A_data.data
        if A_data.dim() > 1:
            A_data = A_data.squeeze(1)                      # (E, 1) -> (E,), executor.py:538-540
        self.ret.data = _flat_spmm(lambda b: spmm(spA.with_weights(A_data), b), B)

    ov._patch(executor.SPMVWithDataExecutor, "run", run_with_data)

    if register_packed:
        # libdgl.so is already loaded RTLD_GLOBAL by the reference (python/dgl/_ffi/base.py:31-36)
        handle = dgl_mod._ffi.base._LIB._handle
        lib = _lib.get_lib()
        n = lib.dglb200_register_packed_funcs(ctypes.c_void_p(handle), int(bool(override_reference)))
        if n < 0:
            raise RuntimeError("dglb200_register_packed_funcs: " + _lib.last_error())
        ov.registered = n
        if override_reference:
            _rebind_globals(ov, dgl_mod)
    return ov


# global names dglb200_register_packed_funcs overrides -> the module attribute that caches each
_OVERRIDDEN = (["runtime.degree_bucketing._CAPI_DGLDegreeBucketing"] +
               ["graph_index._CAPI_DGLGraph" + n for n in
                ("GetAdj", "InEdges_2", "OutEdges_2", "Edges", "InDegrees", "OutDegrees", "FindEdges", "EdgeIds", "InEdges_1",
                 "OutEdges_1", "InDegree", "OutDegree", "AddVertices", "AddEdge",
                 "AddEdges", "Clear", "Free")] +
               ["graph_index._CAPI_DGLGraphUniformSampling" + k for k in ("", "2", "4", "8", "16", "32", "64", "128")])


def _rebind_globals(ov, dgl_mod):
    """Point the reference's cached function objects at the overridden registry entries."""
    import sys
    get = dgl_mod._ffi.function.get_global_func
    for full in _OVERRIDDEN:
        modname, attr = full.rsplit(".", 1)
        mod = sys.modules.get(dgl_mod.__name__ + "." + modname)
        if mod is None or not hasattr(mod, attr):
            continue
        fresh = get(full)
        fresh.is_global = True
        ov._patch(mod, attr, fresh)
    gi = dgl_mod.graph_index
    if hasattr(gi, "_NEIGHBOR_SAMPLING_APIS"):
        table = {k: getattr(gi, "_CAPI_DGLGraphUniformSampling" + ("" if k == 1 else str(k)))
                 for k in gi._NEIGHBOR_SAMPLING_APIS}
        ov._patch(gi, "_NEIGHBOR_SAMPLING_APIS", table)
