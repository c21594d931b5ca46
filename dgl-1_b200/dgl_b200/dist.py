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
import os

import torch
import torch.distributed as dist

from . import ops


# CTAs of a peer pull (max_ctas of the gather kernels; 0 = full grid).  Measured on C5: a pull needs
# the whole GPU's bytes in flight to fill the NVLink (2 ranks: 96 CTAs 290 GB/s, 192: 545, 384 and
# full grid: 660 GB/s; 8 ranks all-to-all: full grid 450 GB/s, 256 CTAs 260 GB/s), so bounding the grid
# to co-schedule the SpMM loses more than the overlap wins: the pipelined paths below (row chunks /
# feature slabs on a side stream) are kept selectable but are OFF by default
PULL_CTAS = int(os.environ.get("DGL_B200_PULL_CTAS", "0"))
# weight of a row relative to an edge in the partition balance (2 SpMMs per step write every row once)
ROW_WEIGHT = float(os.environ.get("DGL_B200_ROW_WEIGHT", "4.0"))
# halo transport over the mapped peer buffers: "perpeer" = one PULL launch per owner in a rotating order
# (every owner serves exactly one puller at a time), "fused" = one pull launch over all owners, "push" =
# the owner gathers and STORES into the requester's buffer (posted NVLink writes)
# measured, C5 x 8 ranks (ms/step): perpeer 14.8, fused 18.4, push 14.5; x 2 ranks: push 26.6, push pipelined
# in 4 row chunks with a 128-CTA push grid 24.7
PULL_MODE = os.environ.get("DGL_B200_PULL", "push")
# precision of the halo rows ON THE WIRE (push transport only): "f32" is the parity mode; "bf16" rounds
# the pushed rows to bfloat16 (half the NVLink bytes: SURVEY.md 8e lever 2), the receiver widens them
# back into its fp32 operand, accumulation stays fp32 - results within the 1e-2 bf16 tolerance.
# Opt-in and NOT yet measured on the GPUs (added after this round's GPU budget was spent).
WIRE = os.environ.get("DGL_B200_WIRE", "f32")
PUSH_CTAS_PIPELINED = 128   # grid of a push that runs beside the SpMM of the previous row chunk
TIMING = bool(int(os.environ.get("DGL_B200_TIMING", "0")))   # record events around the halo transfers
COMM_PRIORITY = -1      # side stream of the pulls: higher priority than the compute stream


def edge_balanced_bounds(indptr_in, indptr_out, parts):
    """Row-block boundaries [0 = b_0 <= ... <= b_parts = n] such that every block holds about
    the same forward + gradient WORK.  Work is the SpMM's algorithmic bytes: every edge gathers one
    feature row and every row writes one, so a row weighs about as much as an
    edge - balancing edges alone leaves ranks with many low-degree rows 40 % slower (measured on
    C5: 11.5 ms vs 16.4 ms per SpMM at 2 ranks)."""
    n = int(indptr_in.shape[0] - 1)
    rows = torch.arange(n + 1, dtype=torch.float64, device=indptr_in.device)
    load = (indptr_in + indptr_out).to(torch.float64) + ROW_WEIGHT * rows
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


def close_peer_mappings():
    """Unmap every peer buffer this process opened with share_with_peers (dglb200_ipc_close).  Call it
    once the tensors returned by share_with_peers are no longer used (e.g. before tearing a
    partition down in a long-lived process); process exit unmaps them anyway."""
    from . import _lib
    lib = _lib.get_lib()
    while _OPEN_HANDLES:
        ptr = _OPEN_HANDLES.pop()
        _lib.check_call(lib.dglb200_ipc_close(ptr))


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
        self._comm = None

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

    def operand(self, key, feat, like):
        """The gather operand as the last exchange left it (benchmark kernel-only timing)."""
        return self.full_buffer(key, feat, like)

    def exchange_bytes(self, feat, elem=4):
        """Bytes this rank receives per exchange."""
        return (self.n - self.rows) * feat * elem if self.world > 1 else 0

    SLAB = 0      # feature-slab pipeline width in elements (64 = the SpMM's L2 slab at C2); 0 = off:
                  # measured at 8 ranks on C2: 2.63 ms/step pipelined vs 2.48 ms/step plain

    def _aggregate(self, key, local, csr, out):
        """Exchange + SpMM.  With the peer transport the feature dimension is pipelined in slabs:
        slab s+1 is copied out of the owners' memory (NVLink) on a side stream while the SpMM runs
        over slab s (dglb200_spmm_csr_strided on column blocks of the gather buffer)."""
        feat = local.shape[1]
        if key not in self._peer_full or self.SLAB <= 0 or feat <= self.SLAB or feat % self.SLAB:
            return self._spmm(csr, self.exchange(key, local), out)
        full = self.full_buffer(key, feat, local)
        views = self.views(full)
        if views[self.rank].data_ptr() != local.data_ptr():
            views[self.rank].copy_(local)
        if out is None:
            out = torch.empty((csr.n_rows, feat), dtype=local.dtype, device=local.device)
        if self._comm is None:
            self._comm = torch.cuda.Stream(device=local.device, priority=COMM_PRIORITY)
        cur = torch.cuda.current_stream(local.device)
        peers = self._peer_full[key]
        self._barrier()                                           # every owner's block is written
        self._comm.wait_stream(cur)
        events = []
        with torch.cuda.stream(self._comm):
            for s0 in range(0, feat, self.SLAB):
                for j in range(1, self.world):
                    q = (self.rank + j) % self.world
                    b0, b1 = self.bounds[q], self.bounds[q + 1]
                    if b1 > b0:
                        ops.gather_rows(peers[q][b0:b1, s0:s0 + self.SLAB], None,
                                        out=full[b0:b1, s0:s0 + self.SLAB], max_ctas=PULL_CTAS)
                ev = torch.cuda.Event()
                ev.record(self._comm)
                events.append(ev)
        for i, s0 in enumerate(range(0, feat, self.SLAB)):
            cur.wait_event(events[i])
            self._spmm(csr, full[:, s0:s0 + self.SLAB], out[:, s0:s0 + self.SLAB])
        self._barrier()                                           # nobody overwrites a block still being read
        return out

    def forward(self, x_local, out=None, key="x"):
        """y_local = (A x)[r0:r1] for update_all(copy_src, sum).  `key` names the gather buffer: a
        model that aggregates at several widths uses one key per width (no reallocation)."""
        return self._aggregate(key, x_local, self.loc_in, out)

    def backward(self, dy_local, out=None, key="dy"):
        """dx_local = (A^T dy)[r0:r1]."""
        return self._aggregate(key, dy_local, self.loc_out, out)


class _PartitionAggregate(torch.autograd.Function):
    """Differentiable update_all(copy_src, sum) over a row partition: forward = exchange + SpMM over
    the in-edge row block, backward = exchange of dy + SpMM over the out-edge row block (what torch
    autograd does for th.sparse.mm in the reference, python/dgl/backend/pytorch/tensor.py:143-146)."""

    @staticmethod
    def forward(ctx, part, x_local, tag):
        ctx.part, ctx.tag = part, tag
        return part.forward(x_local.contiguous(), key="x" + tag)

    @staticmethod
    def backward(ctx, dy):
        return None, ctx.part.backward(dy.contiguous(), key="dy" + ctx.tag), None


def aggregate(part, x_local, tag=""):
    """update_all(copy_src, sum) on this rank's rows of a RowPartition, with autograd."""
    return _PartitionAggregate.apply(part, x_local, tag)


class HaloExchange(object):
    """Compacted, pipelined halo plan for one CSR row block (a full all-gather of X
    moves n * feat per rank; only the DISTINCT remote rows a rank's edges reference are needed -
    measured on C5 at 8 ranks: 9 % of the full volume, 2.0 GB instead of 23.5 GB per rank).

    Plan time (one-off per graph and partition, integer plumbing):
      * the block's rows are cut into `chunks` edge-balanced row chunks; chunk k's halo is the set
        of remote column ids it references that no earlier chunk referenced
      * the compact row space is [own rows | halo of chunk 0 (by owner, ascending id) | chunk 1 ..]
        and the block's column ids are remapped into it
      * an all-to-all of the id lists tells every owner which rows to serve, per chunk
    Per call, for k = 0..chunks-1 on a side stream: fetch chunk k's halo rows (NCCL: pack with the
    gather kernel + all_to_all_single; peer transport: the gather kernel pulls them straight out of
    the owners' memory over NVLink); the SpMM over chunk k's rows starts as soon as its halo has
    landed, while chunk k+1's halo is in flight.
    """

    def __init__(self, csr_block, bounds, rank, world, group=None, gather_fn=None, chunks=1):
        self.group, self.rank, self.world = group, rank, world
        self.bounds = bounds
        self.r0, self.r1 = bounds[rank], bounds[rank + 1]
        self.rows = self.r1 - self.r0
        self._gather = gather_fn or ops.gather_rows
        dev = csr_block.indices.device
        self.chunks = max(1, int(chunks)) if world > 1 else 1
        # edge-balanced row chunks of the block
        ip = csr_block.indptr
        tgt = torch.arange(1, self.chunks, dtype=torch.float64, device=dev) * (float(ip[-1]) / self.chunks)
        cuts = torch.searchsorted(ip.to(torch.float64), tgt).clamp_(max=self.rows).tolist()
        self.row_cuts = [0] + [int(c) for c in cuts] + [self.rows]
        for i in range(1, len(self.row_cuts)):
            self.row_cuts[i] = max(self.row_cuts[i], self.row_cuts[i - 1])
        cols = csr_block.indices.long()
        own = (cols >= self.r0) & (cols < self.r1)
        halo_ids = torch.unique(cols[~own])                       # H: sorted ascending, own range absent
        # first chunk that references every halo id
        first = torch.full((halo_ids.numel(),), self.chunks, dtype=torch.int64, device=dev)
        for k in range(self.chunks):
            lo, hi = int(ip[self.row_cuts[k]]), int(ip[self.row_cuts[k + 1]])
            ck = cols[lo:hi]
            ck = ck[(ck < self.r0) | (ck >= self.r1)]
            hk = torch.searchsorted(halo_ids, torch.unique(ck))
            first[hk] = torch.minimum(first[hk], torch.full_like(hk, k))
            del ck, hk
        order = torch.sort(first, stable=True)[1]                 # layout: (chunk, id) == (chunk, owner, id)
        pos = torch.empty_like(order)
        pos[order] = torch.arange(order.numel(), device=dev)
        layout_ids = halo_ids[order]                              # global ids in layout order
        per_chunk = torch.bincount(first, minlength=self.chunks).tolist()
        self.n_halo = int(halo_ids.numel())
        new_cols = torch.where(own, cols - self.r0,
                               self.rows + pos[torch.searchsorted(halo_ids, cols).clamp_(max=max(self.n_halo - 1, 0))]
                               if self.n_halo else cols - self.r0)
        self.csr = ops.CSR(csr_block.n_rows, self.rows + self.n_halo, csr_block.indptr,
                           new_cols.to(torch.int32), csr_block.eids)
        del cols, own, new_cols, pos, order, first
        # per (chunk, owner): receive counts, and the ids as LOCAL row ids of their owner
        b = torch.tensor(bounds, dtype=torch.int64, device=dev)
        self.chunk_start, self.recv_counts, self.need_local = [], [], []
        off = 0
        for k in range(self.chunks):
            ids_k = layout_ids[off:off + per_chunk[k]]
            cut = torch.searchsorted(ids_k, b)
            cnt = (cut[1:] - cut[:-1]).tolist()
            self.chunk_start.append(self.rows + off)
            self.recv_counts.append(cnt)
            owner_of = torch.repeat_interleave(torch.arange(world, device=dev), torch.tensor(cnt, device=dev))
            local_k = ids_k - b[owner_of]
            self.need_local.append([local_k[int(cut[q]):int(cut[q + 1])].contiguous() for q in range(world)])
            off += per_chunk[k]
        # tell every owner which of its rows I need, per chunk
        self.send_counts, self.serve_local, self.need_cat = [], [], []
        for k in range(self.chunks):
            rc = torch.tensor(self.recv_counts[k], dtype=torch.int64, device=dev)
            sc = torch.empty_like(rc)
            if world > 1:
                dist.all_to_all_single(sc, rc, group=group)
            else:
                sc.copy_(rc)
            sc = sc.tolist()
            need_cat = torch.cat(self.need_local[k]) if world > 1 else torch.empty(0, dtype=torch.int64, device=dev)
            serve = torch.empty(int(sum(sc)), dtype=torch.int64, device=dev)
            if world > 1:
                dist.all_to_all_single(serve, need_cat, output_split_sizes=sc,
                                       input_split_sizes=self.recv_counts[k], group=group)
            self.send_counts.append(sc)
            self.need_cat.append(need_cat.contiguous())
            self.serve_local.append(serve.contiguous())           # already local ids of THIS rank
        # push transport: where the rows I serve land inside requester r's compact buffer, per chunk
        self.push_dst = None
        if world > 1:
            objs = [None] * world
            dist.all_gather_object(objs, (self.chunk_start, self.recv_counts), group=group)
            self.push_dst = [[int(objs[r][0][k] + sum(objs[r][1][k][:rank])) for r in range(world)]
                             for k in range(self.chunks)]
        self.chunk_csr = None
        self.timing = []
        self._buf = None
        self._peer_buf = None
        self._wire = self._peer_wire = None     # bf16 wire mode (enable_peer_transport)
        self._barrier = None
        self._comm = None

    # -- buffers / transports -----------------------------------------------------------------
    def buffer(self, feat, like):
        """Compact feature buffer [rows + n_halo, feat]; rows [0, rows) are this rank's block."""
        if (self._buf is None or self._buf.shape[1] != feat or self._buf.dtype != like.dtype or
                self._buf.device != like.device):
            self._buf = torch.empty((self.rows + self.n_halo, feat), dtype=like.dtype, device=like.device)
        return self._buf

    def enable_peer_transport(self, feat, like):
        """Pull the halo rows straight out of their owners' memory with the gather kernel (peer
        loads over NVLink): no packing pass on the owner, no NCCL on the data path."""
        if self.world == 1:
            return
        buf = self.buffer(feat, like)
        self._peer_buf = share_with_peers(buf, self.rank, self.world, self.group)
        self._barrier = StreamBarrier(like.device, self.group)
        self._wire = self._peer_wire = None
        if WIRE == "bf16" and PULL_MODE == "push" and buf.dtype == torch.float32:
            # halo rows arrive here as bf16 (written by their owners) and are widened into `buf`
            self._wire = torch.empty((max(self.n_halo, 1), feat), dtype=torch.bfloat16, device=like.device)
            self._peer_wire = share_with_peers(self._wire, self.rank, self.world, self.group)
            rows_all = [None] * self.world
            dist.all_gather_object(rows_all, self.rows, group=self.group)
            self._rows_of_rank = rows_all

    def row_chunks(self):
        """The remapped CSR cut into the plan's row chunks (views; each with its own long-row plan)."""
        if self.chunk_csr is None:
            self.chunk_csr = [ops.csr_row_block(self.csr, self.row_cuts[k], self.row_cuts[k + 1])
                              for k in range(self.chunks)]
        return self.chunk_csr

    def halo_bytes(self, feat, elem=4):
        return self.n_halo * feat * elem

    # -- data path ---------------------------------------------------------------------------
    def _fetch_chunk(self, k, buf):
        """Bring chunk k's halo rows into buf (current stream)."""
        start = self.chunk_start[k]
        cnts = self.recv_counts[k]
        if self._peer_buf is not None and PULL_MODE == "push":
            # owner-side: gather the rows each requester needs out of my block and STORE them straight
            # into its compact buffer over NVLink (posted writes, no request round trip); rotating
            # order so that every requester is written by one owner at a time
            sc = self.send_counts[k]
            seg = [0]
            for r in range(self.world):
                seg.append(seg[-1] + sc[r])
            ctas = PULL_CTAS or (PUSH_CTAS_PIPELINED if self.chunks > 1 else 0)
            for j in range(1, self.world):
                r = (self.rank + j) % self.world
                if sc[r]:
                    d0 = self.push_dst[k][r]
                    if self._peer_wire is not None:
                        h0 = d0 - self._rows_of_rank[r]          # position inside r's halo region
                        ops.gather_rows_convert(buf[:self.rows], self.serve_local[k][seg[r]:seg[r + 1]],
                                                out=self._peer_wire[r][h0:h0 + sc[r]], max_ctas=ctas)
                    else:
                        ops.gather_rows(buf[:self.rows], self.serve_local[k][seg[r]:seg[r + 1]],
                                        out=self._peer_buf[r][d0:d0 + sc[r]], max_ctas=ctas)
            return
        if self._peer_buf is not None:
            # one launch pulls this chunk's rows from every owner at once (NVLink peer loads)
            seg = [0]
            for q in range(self.world):
                seg.append(seg[-1] + cnts[q])
            if seg[-1] and PULL_MODE == "perpeer":
                for j in range(1, self.world):
                    q = (self.rank + j) % self.world
                    if cnts[q]:
                        ops.gather_rows(self._peer_buf[q][:self.bounds[q + 1] - self.bounds[q]], self.need_local[k][q],
                                        out=buf[start + seg[q]:start + seg[q + 1]], max_ctas=PULL_CTAS)
            elif seg[-1]:
                srcs = [self._peer_buf[q][:self.bounds[q + 1] - self.bounds[q]] if q != self.rank else None
                        for q in range(self.world)]
                ops.gather_rows_multi(srcs, seg, self.need_cat[k], buf[start:start + seg[-1]], max_ctas=PULL_CTAS)
            return
        send = self._gather(buf[:self.rows], self.serve_local[k])
        dist.all_to_all_single(buf[start:start + sum(cnts)], send, output_split_sizes=cnts,
                               input_split_sizes=self.send_counts[k], group=self.group)

    def _widen_chunk(self, k, buf):
        """bf16 wire mode: chunk k's halo rows have landed in the bf16 staging buffer; widen them
        into the fp32 operand (local HBM traffic only)."""
        if self._wire is None:
            return
        cnt = sum(self.recv_counts[k])
        if cnt:
            h0 = self.chunk_start[k] - self.rows
            ops.gather_rows_convert(self._wire[h0:h0 + cnt], None,
                                    out=buf[self.chunk_start[k]:self.chunk_start[k] + cnt])

    def exchange(self, x_local):
        """Unpipelined: returns the compact buffer with own rows and every halo row."""
        buf = self.buffer(x_local.shape[1], x_local)
        own = buf[:self.rows]
        if own.data_ptr() != x_local.data_ptr():
            own.copy_(x_local)
        if self.world == 1:
            return buf
        if self._barrier is not None:
            self._barrier()
        if TIMING and x_local.is_cuda:
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
        for k in range(self.chunks):
            self._fetch_chunk(k, buf)
        if TIMING and x_local.is_cuda:
            b.record()
            self.timing.append((a, b))
        if self._barrier is not None:
            self._barrier()
        for k in range(self.chunks):
            self._widen_chunk(k, buf)
        return buf

    def aggregate(self, x_local, spmm_fn, out):
        """Pipelined exchange + SpMM: y[rows of chunk k] is computed while chunk k+1's halo is in
        flight on the side stream.  `out` [rows, feat] must be given."""
        if self.world == 1 or self.chunks == 1:
            return spmm_fn(self.csr, self.exchange(x_local), out)
        if self._peer_buf is not None and PULL_MODE == "push":
            return self._aggregate_pushed(x_local, spmm_fn, out)
        buf = self.buffer(x_local.shape[1], x_local)
        own = buf[:self.rows]
        if own.data_ptr() != x_local.data_ptr():
            own.copy_(x_local)
        if not x_local.is_cuda:                                   # CPU (gloo tests): same schedule, no streams
            for k, csr_k in enumerate(self.row_chunks()):
                self._fetch_chunk(k, buf)
                if csr_k.n_rows:
                    out[self.row_cuts[k]:self.row_cuts[k + 1]] = spmm_fn(csr_k, buf, None)
            return out
        if self._comm is None:
            # high priority + bounded grid: the pull's few CTAs are always dispatched at once and
            # the SpMM of the previous chunk fills the remaining SM slots
            self._comm = torch.cuda.Stream(device=x_local.device, priority=COMM_PRIORITY)
        cur = torch.cuda.current_stream(x_local.device)
        if self._barrier is not None:
            self._barrier()                                       # every owner's rows are written
        self._comm.wait_stream(cur)
        events = []
        with torch.cuda.stream(self._comm):
            for k in range(self.chunks):
                self._fetch_chunk(k, buf)
                ev = torch.cuda.Event()
                ev.record(self._comm)
                events.append(ev)
        for k, csr_k in enumerate(self.row_chunks()):
            cur.wait_event(events[k])
            if csr_k.n_rows:
                spmm_fn(csr_k, buf, out[self.row_cuts[k]:self.row_cuts[k + 1]])
        if self._barrier is not None:
            self._barrier()                                       # nobody overwrites rows still being read
        return out


    def _aggregate_pushed(self, x_local, spmm_fn, out):
        """Pipelined PUSH exchange + SpMM.  A push is local loads + posted NVLink stores, so (unlike
        a pull, which needs the whole GPU's loads in flight to cover the NVLink round trip) a small
        grid keeps the link busy and the SpMM of the previous row chunk keeps the other SMs.  A
        pushed chunk is complete on the receiver only when EVERY owner has stored its share: one
        stream-ordered barrier per chunk on the side stream."""
        buf = self.buffer(x_local.shape[1], x_local)
        own = buf[:self.rows]
        if own.data_ptr() != x_local.data_ptr():
            own.copy_(x_local)
        if self._comm is None:
            self._comm = torch.cuda.Stream(device=x_local.device, priority=COMM_PRIORITY)
        cur = torch.cuda.current_stream(x_local.device)
        self._barrier()                       # every rank's previous SpMM has finished reading its buffer
        self._comm.wait_stream(cur)
        events = []
        with torch.cuda.stream(self._comm):
            for k in range(self.chunks):
                if TIMING:
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record()
                self._fetch_chunk(k, buf)
                if TIMING:
                    b.record()
                    self.timing.append((a, b))
                self._barrier()               # chunk k's halo has landed everywhere
                self._widen_chunk(k, buf)
                ev = torch.cuda.Event()
                ev.record(self._comm)
                events.append(ev)
        for k, csr_k in enumerate(self.row_chunks()):
            cur.wait_event(events[k])
            if csr_k.n_rows:
                spmm_fn(csr_k, buf, out[self.row_cuts[k]:self.row_cuts[k + 1]])
        return out


class HaloPartition(RowPartition):
    """RowPartition whose exchange is the compacted (and, with chunks > 1, pipelined) halo instead
    of the full all-gather."""

    def __init__(self, csr_in, csr_out, rank=None, world=None, group=None, spmm_fn=None, gather_fn=None,
                 chunks=1):
        super(HaloPartition, self).__init__(csr_in, csr_out, rank, world, group, spmm_fn)
        self.halo_in = HaloExchange(self.loc_in, self.bounds, self.rank, self.world, group, gather_fn, chunks)
        self.halo_out = HaloExchange(self.loc_out, self.bounds, self.rank, self.world, group, gather_fn, chunks)
        self.loc_in, self.loc_out = self.halo_in.csr, self.halo_out.csr

    def _halo(self, key):
        return self.halo_in if key.startswith("x") else self.halo_out

    def enable_peer_transport(self, feat, like, keys=("x", "dy")):
        for key in keys:
            self._halo(key).enable_peer_transport(feat, like)

    def local_view(self, key, feat, like):
        return self._halo(key).buffer(feat, like)[:self.rows]

    def exchange(self, key, local):
        return self._halo(key).exchange(local)

    def operand(self, key, feat, like):
        return self._halo(key).buffer(feat, like)

    def exchange_bytes(self, feat, elem=4):
        if self.halo_in._wire is not None:
            elem = 2
        return max(self.halo_in.n_halo, self.halo_out.n_halo) * feat * elem

    def forward(self, x_local, out=None, key="x"):
        if out is None:
            out = torch.empty((self.rows, x_local.shape[1]), dtype=x_local.dtype, device=x_local.device)
        return self.halo_in.aggregate(x_local, self._spmm, out)

    def backward(self, dy_local, out=None, key="dy"):
        if out is None:
            out = torch.empty((self.rows, dy_local.shape[1]), dtype=dy_local.dtype, device=dy_local.device)
        return self.halo_out.aggregate(dy_local, self._spmm, out)
