"""Multi-GPU message passing: 1-D destination-row partition + feature-block all-gather.

The reference has no distributed layer; this is the north-star's
scheme: one process per GPU, contiguous edge-balanced row blocks, rank p owns the rows
[bounds[p], bounds[p+1]) of every node tensor and of both CSRs.  Forward: all-gather the source
feature blocks (the halo exchange), then the local SpMM over the rank's destination rows.
Gradient: all-gather the dY blocks, then the local transposed SpMM over the rank's source rows
(gather form: no reduce-scatter, no atomics, deterministic).

Only torch.distributed is used for the exchange (NCCL on GPUs, gloo in the CPU tests); the
local compute is `ops.spmm` (libdgl_b200.so) unless a test injects a checker.
"""
import torch
import torch.distributed as dist

from . import ops


def edge_balanced_bounds(indptr_in, indptr_out, parts):
    """Row-block boundaries [0 = b_0 <= ... <= b_parts = n] such that every block holds about
    the same number of in + out edges (forward and gradient work)."""
    load = (indptr_in + indptr_out).to(torch.float64)
    n = int(indptr_in.shape[0] - 1)
    total = float(load[-1])
    if parts == 1 or total == 0:
        step = (n + parts - 1) // parts
        return [min(n, p * step) for p in range(parts)] + [n]
    targets = torch.arange(1, parts, dtype=torch.float64, device=load.device) * (total / parts)
    cuts = torch.searchsorted(load, targets).clamp_(max=n).tolist()
    bounds = [0] + [int(c) for c in cuts] + [n]
    for i in range(1, len(bounds)):            # monotone
        bounds[i] = max(bounds[i], bounds[i - 1])
    return bounds


class _RawCuda(object):
    """__cuda_array_interface__ view of a raw device pointer (zero copy)."""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr,
                                         "data": (int(ptr), False), "version": 2}


_TYPESTR = {torch.float32: ("<f4", None), torch.bfloat16: ("<u2", torch.bfloat16),
            torch.int64: ("<i8", None), torch.int32: ("<i4", None)}
_OPEN_HANDLES = []


def share_with_peers(t, rank, world, group=None):
    """Map every rank's contiguous CUDA tensor `t` into every other rank's address space with CUDA
    IPC (dglb200_ipc_get_handle / dglb200_ipc_open: the mapping is created on the IMPORTING rank's
    own device, which also enables peer access).  Returns `world` tensors; entry q aliases GPU q's
    memory and can be read by this rank's kernels over NVLink (measured: random 512-byte rows
    pulled at 670 GB/s by dglb200_gather_rows).  The owner must keep `t` alive."""
    import ctypes
    from . import _lib
    lib = _lib.get_lib()
    assert t.is_cuda and t.is_contiguous()
    handle = ctypes.create_string_buffer(64)
    offset = ctypes.c_int64(0)
    with torch.cuda.device(t.device):
        _lib.check_call(lib.dglb200_ipc_get_handle(ctypes.c_void_p(t.data_ptr()), handle, ctypes.byref(offset)))
    objs = [None] * world
    dist.all_gather_object(objs, (handle.raw, int(offset.value), tuple(t.shape), str(t.dtype)), group=group)
    typestr, view_as = _TYPESTR[t.dtype]
    peers = []
    for q in range(world):
        if q == rank:
            peers.append(t)
            continue
        h, off, shape, dtype = objs[q]
        assert dtype == str(t.dtype)
        ptr = ctypes.c_void_p()
        with torch.cuda.device(t.device):
            _lib.check_call(lib.dglb200_ipc_open(ctypes.c_char_p(h), ctypes.byref(ptr)))
        _OPEN_HANDLES.append(ptr)
        pt = torch.as_tensor(_RawCuda(ptr.value + off, shape, typestr), device=t.device)
        peers.append(pt.view(view_as) if view_as is not None else pt)
    return peers


class StreamBarrier(object):
    """Cross-rank barrier in STREAM order (no host stall): a 1-element all-reduce.  Every rank's
    kernels enqueued before it are complete before any rank's kernels enqueued after it start."""

    def __init__(self, device, group=None):
        self.token = torch.zeros(1, dtype=torch.float32, device=device)
        self.group = group

    def __call__(self):
        dist.all_reduce(self.token, group=self.group)


class RowPartition(object):
    """This rank's share of a graph: row blocks of the in-edge CSR (forward) and of the out-edge
    CSR (gradient), plus the all-gather plumbing for node features."""

    def __init__(self, csr_in, csr_out, rank=None, world=None, group=None, spmm_fn=None):
        self.group = group
        self.world = world if world is not None else (dist.get_world_size(group) if dist.is_initialized() else 1)
        self.rank = rank if rank is not None else (dist.get_rank(group) if dist.is_initialized() else 0)
        self.n = csr_in.n_rows
        self.bounds = edge_balanced_bounds(csr_in.indptr, csr_out.indptr, self.world)
        self.r0, self.r1 = self.bounds[self.rank], self.bounds[self.rank + 1]
        self.rows = self.r1 - self.r0
        if self.world > 1:
            self.loc_in = ops.csr_row_block(csr_in, self.r0, self.r1)
            self.loc_out = ops.csr_row_block(csr_out, self.r0, self.r1)
        else:
            self.loc_in, self.loc_out = csr_in, csr_out
        self._spmm = spmm_fn or (lambda csr, x, out: ops.spmm(csr, ops.MSG_COPY_SRC, ops.RED_SUM, x=x, out=out)[0])
        self._full = {}
        self._peer_full = {}      # key -> every rank's full buffer mapped here (transport "peer")
        self._barrier = None

    def enable_peer_transport(self, feat, like, keys=("x", "dy")):
        """Exchange by PEER LOADS over NVLink instead of NCCL: every rank's gather buffer is mapped
        into every other rank; a rank copies the blocks it lacks straight out of their owners'
        memory (one device-to-device copy per owner, all inside stream-ordered barriers)."""
        if self.world == 1:
            return
        for key in keys:
            full = self.full_buffer(key, feat, like)
            self._peer_full[key] = share_with_peers(full, self.rank, self.world, self.group)
        self._barrier = StreamBarrier(like.device, self.group)

    def _global_rank(self, p):
        return p if self.group is None else dist.get_global_rank(self.group, p)

    def full_buffer(self, key, feat, like):
        """Persistent [n, feat] gather buffer (one per use: 'x' for forward, 'dy' for gradient)."""
        buf = self._full.get(key)
        if buf is None or buf.shape[1] != feat or buf.dtype != like.dtype or buf.device != like.device:
            buf = torch.empty((self.n, feat), dtype=like.dtype, device=like.device)
            self._full[key] = buf
        return buf

    def views(self, full):
        return [full[self.bounds[p]:self.bounds[p + 1]] for p in range(self.world)]

    def all_gather_rows(self, local, key):
        """[rows, feat] block of every rank -> [n, feat] on every rank (uneven block sizes)."""
        full = self.full_buffer(key, local.shape[1], local)
        views = self.views(full)
        if self.world == 1:
            if views[0].data_ptr() != local.data_ptr():
                views[0].copy_(local)
            return full
        if views[self.rank].data_ptr() != local.data_ptr():
            views[self.rank].copy_(local)
        if key in self._peer_full:
            self._barrier()                                   # every owner's block is written
            for k in range(1, self.world):                    # start at a different owner on every rank
                q = (self.rank + k) % self.world
                if views[q].numel():
                    views[q].copy_(self._peer_full[key][q][self.bounds[q]:self.bounds[q + 1]], non_blocking=True)
            self._barrier()                                   # nobody overwrites a block still being read
            return full
        if dist.get_backend(self.group) == "nccl":
            # uneven blocks: NCCL runs them as ONE grouped launch of per-owner broadcasts, all links
            # busy in both directions at once
            dist.all_gather(views, views[self.rank], group=self.group)
            return full
        # gloo (CPU tests) has no uneven all_gather: one broadcast per owner
        works = [dist.broadcast(views[p], src=self._global_rank(p), group=self.group, async_op=True)
                 for p in range(self.world) if views[p].numel() > 0]
        for w in works:
            w.wait()
        return full

    # -- exchange interface (same on HaloPartition) --------------------------------------------
    def local_view(self, key, feat, like):
        """Where this rank's block of the 'x' / 'dy' operand lives inside the exchange buffer:
        filling it in place saves the copy."""
        return self.full_buffer(key, feat, like)[self.r0:self.r1]

    def exchange(self, key, local):
        """Make every source row the local SpMM needs available; returns the gather operand."""
        return self.all_gather_rows(local, key)

    def exchange_bytes(self, feat, elem=4):
        """Bytes this rank receives per exchange."""
        return (self.n - self.rows) * feat * elem if self.world > 1 else 0

    def forward(self, x_local, out=None):
        """y_local = (A x)[r0:r1] for update_all(copy_src, sum)."""
        return self._spmm(self.loc_in, self.exchange("x", x_local), out)

    def backward(self, dy_local, out=None):
        """dx_local = (A^T dy)[r0:r1]."""
        return self._spmm(self.loc_out, self.exchange("dy", dy_local), out)


class HaloExchange(object):
    """Compacted halo plan for one CSR row block (a full all-gather of X moves
    n * feat per rank; only the DISTINCT remote rows a rank's edges reference are needed -
    measured on C5 at 8 ranks: 9 % of the full volume, 2.0 GB instead of 23.5 GB per rank).

    Plan time (one-off per graph and partition, integer plumbing):
      * the sorted unique remote column ids of the block, split by owner       -> what I receive
      * an all-to-all of those id lists                                         -> what I send
      * the block's column ids remapped into the compact row space
        [own rows | halo rows from owner 0 | owner 1 | ...]
    Per call: pack the rows the others need (gather_rows kernel), one all_to_all_single straight
    into the halo region of the compact feature buffer, then the local SpMM over the remapped CSR.
    """

    def __init__(self, csr_block, bounds, rank, world, group=None, gather_fn=None):
        self.group, self.rank, self.world = group, rank, world
        self.bounds = bounds
        self.r0, self.r1 = bounds[rank], bounds[rank + 1]
        self.rows = self.r1 - self.r0
        self._gather = gather_fn or ops.gather_rows
        dev = csr_block.indices.device
        cols = csr_block.indices.long()
        own = (cols >= self.r0) & (cols < self.r1)
        halo_ids = torch.unique(cols[~own])                       # sorted ascending, own range absent
        b = torch.tensor(bounds, dtype=torch.int64, device=dev)
        cut = torch.searchsorted(halo_ids, b)
        self.recv_counts = (cut[1:] - cut[:-1]).tolist()          # rows I need from every owner
        assert self.recv_counts[rank] == 0
        new_cols = torch.where(own, cols - self.r0, self.rows + torch.searchsorted(halo_ids, cols))
        self.n_halo = int(halo_ids.numel())
        self.csr = ops.CSR(csr_block.n_rows, self.rows + self.n_halo, csr_block.indptr,
                           new_cols.to(torch.int32), csr_block.eids)
        del cols, own, new_cols
        # tell every owner which of its rows I need
        rc = torch.tensor(self.recv_counts, dtype=torch.int64, device=dev)
        sc = torch.empty_like(rc)
        if world > 1:
            dist.all_to_all_single(sc, rc, group=group)
        else:
            sc.copy_(rc)
        self.send_counts = sc.tolist()
        serve = torch.empty(int(sum(self.send_counts)), dtype=torch.int64, device=dev)
        if world > 1:
            dist.all_to_all_single(serve, halo_ids, output_split_sizes=self.send_counts,
                                   input_split_sizes=self.recv_counts, group=group)
        self.serve_local = (serve - self.r0).contiguous()        # my rows to pack, grouped by requester
        self._buf = None
        self._peer_buf = None
        self._barrier = None
        # for the peer transport: the rows I need, as LOCAL row ids of their owners, per owner
        self.need_local = []
        off = 0
        for q in range(world):
            ids = halo_ids[off:off + self.recv_counts[q]] - bounds[q]
            self.need_local.append(ids.contiguous())
            off += self.recv_counts[q]

    def enable_peer_transport(self, feat, like):
        """Pull the halo rows straight out of their owners' memory with the gather kernel (peer
        loads over NVLink): no packing pass on the owner, no NCCL on the data path."""
        if self.world == 1:
            return
        buf = self.buffer(feat, like)
        self._peer_buf = share_with_peers(buf, self.rank, self.world, self.group)
        self._barrier = StreamBarrier(like.device, self.group)

    def buffer(self, feat, like):
        """Compact feature buffer [rows + n_halo, feat]; rows [0, rows) are this rank's block."""
        if (self._buf is None or self._buf.shape[1] != feat or self._buf.dtype != like.dtype or
                self._buf.device != like.device):
            self._buf = torch.empty((self.rows + self.n_halo, feat), dtype=like.dtype, device=like.device)
        return self._buf

    def exchange(self, x_local):
        """Returns the compact buffer with own rows and freshly received halo rows."""
        buf = self.buffer(x_local.shape[1], x_local)
        own = buf[:self.rows]
        if own.data_ptr() != x_local.data_ptr():
            own.copy_(x_local)
        if self._peer_buf is not None:
            self._barrier()
            for k in range(1, self.world):
                q = (self.rank + k) % self.world
                start = self.rows + sum(self.recv_counts[:q])
                cnt = self.recv_counts[q]
                if cnt:
                    rows_q = self.bounds[q + 1] - self.bounds[q]
                    self._gather(self._peer_buf[q][:rows_q], self.need_local[q], out=buf[start:start + cnt])
            self._barrier()
            return buf
        if self.world > 1:
            send = self._gather(own, self.serve_local)
            dist.all_to_all_single(buf[self.rows:], send, output_split_sizes=self.recv_counts,
                                   input_split_sizes=self.send_counts, group=self.group)
        return buf

    def halo_bytes(self, feat, elem=4):
        return self.n_halo * feat * elem


class HaloPartition(RowPartition):
    """RowPartition whose exchange is the compacted halo instead of the full all-gather."""

    def __init__(self, csr_in, csr_out, rank=None, world=None, group=None, spmm_fn=None, gather_fn=None):
        super(HaloPartition, self).__init__(csr_in, csr_out, rank, world, group, spmm_fn)
        self.halo_in = HaloExchange(self.loc_in, self.bounds, self.rank, self.world, group, gather_fn)
        self.halo_out = HaloExchange(self.loc_out, self.bounds, self.rank, self.world, group, gather_fn)
        self.loc_in, self.loc_out = self.halo_in.csr, self.halo_out.csr

    def _halo(self, key):
        return self.halo_in if key == "x" else self.halo_out

    def enable_peer_transport(self, feat, like, keys=("x", "dy")):
        for key in keys:
            self._halo(key).enable_peer_transport(feat, like)

    def local_view(self, key, feat, like):
        return self._halo(key).buffer(feat, like)[:self.rows]

    def exchange(self, key, local):
        return self._halo(key).exchange(local)

    def exchange_bytes(self, feat, elem=4):
        return max(self.halo_in.n_halo, self.halo_out.n_halo) * feat * elem
