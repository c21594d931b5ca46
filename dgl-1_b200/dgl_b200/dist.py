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
from .base import DGLError


# CTAs of a peer pull (max_ctas of the gather kernels; 0 = full grid).  Measured on C5: a pull needs
# the whole GPU's bytes in flight to fill the NVLink (2 ranks: 96 CTAs 290 GB/s, 192: 545, 384 and
# full grid: 660 GB/s; 8 ranks all-to-all: full grid 450 GB/s, 256 CTAs 260 GB/s), so bounding the grid
# to co-schedule the SpMM loses more than the overlap wins: the pipelined paths below (row chunks /
# feature slabs on a side stream) are kept selectable but are OFF by default
PULL_CTAS = int(os.environ.get("DGL_B200_PULL_CTAS", "0"))
# weight of a row relative to an edge in the partition balance (2 SpMMs per step write every row once)
ROW_WEIGHT = float(os.environ.get("DGL_B200_ROW_WEIGHT", "4.0"))
# Transport over the mapped peer buffers:
#   "signal"  (default) ONE push launch per chunk stores this rank's rows into EVERY peer's buffer over
#             NVLink and then sets a flag word in each peer's memory; the consumer's stream waits on its
#             own flag words (dglb200_peer_push / peer_wait / peer_signal).  No NCCL call, no stream
#             barrier and no host round trip on the data path; pushes of chunk k+1 run on a high-priority
#             side stream beside the SpMM of chunk k.
#   "push"    round 1: the owner gathers and STORES into the requester's buffer, one launch per peer in
#             a rotating order, chunks separated by stream-ordered NCCL barriers (1-element all-reduce)
#   "perpeer" one PULL launch per owner in a rotating order; "fused": one pull launch over all owners
# measured, C5 x 8 ranks (ms/step, round 1): perpeer 14.8, fused 18.4, push 14.5, push pipelined 11.3
PULL_MODE = os.environ.get("DGL_B200_PULL", "signal")
# precision of the halo rows ON THE WIRE (push transports only): "f32" is the parity mode; "bf16" rounds
# the pushed rows to bfloat16 (half the NVLink bytes: SURVEY.md 8e lever 2), the receiver widens them
# back into its fp32 operand, accumulation stays fp32 - results within the 1e-2 bf16 tolerance.
WIRE = os.environ.get("DGL_B200_WIRE", "f32")
PUSH_CTAS_PIPELINED = 128   # "push" mode: grid of a push that runs beside the SpMM of the previous row chunk
PUSH_CTAS = int(os.environ.get("DGL_B200_PUSH_CTAS", "0"))   # "signal" mode: grid bound of a push (0 = 2 CTAs per SM)
TIMING = bool(int(os.environ.get("DGL_B200_TIMING", "0")))   # record events around the halo transfers
COMM_PRIORITY = -1      # side stream of the exchange: higher priority than the compute stream
# bf16 wire: the SpMM can gather the halo rows straight from the bf16 staging buffer (dglb200_spmm_csr_ext,
# mixed-precision operand) instead of widening them into the fp32 buffer first.  OFF by default - measured
# on a real C5 rank block (8-way, 6.0M rows + 4.0M halo rows, 19.2M edges): 4.0 ms against 2.5 ms for the
# fp32 gather (the widening adds ~10 instructions per gathered 16 bytes to a loop that already runs at 59 %
# issue utilisation, and at 64 registers it spills); the separate widening pass costs 0.6 ms per op
MIXED_HALO = bool(int(os.environ.get("DGL_B200_MIXED_HALO", "0")))
# all-gather partition: sum the column block of the rank's OWN source rows first (needs no exchange, so it
# runs while every slab is still in flight), then add the remote column block slab by slab
SPLIT_LOCAL = bool(int(os.environ.get("DGL_B200_SPLIT_LOCAL", "1")))
# feature-slab width (elements) of the pipelined all-gather; 0 = by rank count: 64 (the SpMM's own L2 slab at C2)
# up to 2 ranks, 128 beyond - measured on C2 x 8 (ms/step): 64 -> 2.49, 128 -> 2.12, 256 (one push) -> 2.16: at 8
# ranks an op is ~1 ms and the per-slab launches (push, wait, SpMM, fix-up) start to bound it
ALLGATHER_SLAB = int(os.environ.get("DGL_B200_AG_SLAB", "0"))
ALLGATHER_SLABS = os.environ.get("DGL_B200_AG_SLABS", "")       # e.g. "64,192": widths that sum to the feature width


def chunk_fractions(chunks):
    """Cumulative edge fractions at which a row block is cut into pipeline chunks.  The exchange is the
    critical path on sparse graphs (C5 x 8: 3.3 ms of pushes against 2.5 ms of SpMM per op), so what is
    left exposed is the first chunk's push and the LAST chunk's SpMM: both are kept small, the middle
    chunks carry the bulk.  DGL_B200_CHUNK_FRACS="0.15,0.4,0.7,0.9,1" overrides."""
    env = os.environ.get("DGL_B200_CHUNK_FRACS")
    if env:
        fr = [float(v) for v in env.split(",")]
        if len(fr) == chunks and abs(fr[-1] - 1.0) < 1e-9 and all(b > a for a, b in zip(fr, fr[1:])):
            return fr
    if chunks <= 2:
        return [(k + 1) / chunks for k in range(chunks)]
    # first and last chunk half the weight of the middle ones
    w = [0.5] + [1.0] * (chunks - 2) + [0.5]
    tot, acc, fr = sum(w), 0.0, []
    for v in w:
        acc += v
        fr.append(acc / tot)
    fr[-1] = 1.0
    return fr


# How a push reaches the peer (csrc/peer.cu): 0 = every thread's st.global (a push CTA then moves only ~5 GB/s,
# the link needs ~300 CTAs and the SpMM sharing the SMs slows down); 1 = threads gather (+ bf16 rounding) into
# shared memory, ONE thread hands each stage to the TMA (cp.async.bulk shared -> peer global); 2 = TMA both ways
# (bulk gathers global -> shared behind an mbarrier, bulk stores to the peer; fp32 wire only, else 1).
# Measured, C5 x 8, 4 chunks, ms/step: fp32 wire 10.79 / 10.44 / 10.39; bf16 wire 9.30 / 9.00 / -
PUSH_TMA = int(os.environ.get("DGL_B200_PUSH_TMA", "2"))


def apply_tuning():
    """Hand the transport knobs that live in the library to it (dglb200_set_tuning)."""
    from . import _lib
    _lib.check_call(_lib.get_lib().dglb200_set_tuning(b"peer_push_tma", int(PUSH_TMA)))


def transport_name(exchange):
    """Human-readable name of the peer transport in use (bench.py's `config.parallelism`)."""
    if PULL_MODE == "signal":
        return "push (%s) + flag signals (own kernels, no NCCL on the data path)" % (
            {0: "thread stores", 1: "TMA bulk stores", 2: "TMA bulk gathers + stores"}.get(PUSH_TMA, "?"))
    return PULL_MODE if exchange == "halo" else "block copies between NCCL stream barriers"


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
_OPEN_HANDLES = {}        # IPC handle bytes -> mapped base pointer (one mapping per exporter allocation)


def share_with_peers(t, rank, world, group=None):
    """Map every rank's contiguous CUDA tensor `t` into every other rank's address space with CUDA
    IPC (dglb200_ipc_get_handle / dglb200_ipc_open: the mapping is created on the IMPORTING rank's
    own device, which also enables peer access).  Returns `world` tensors; entry q aliases GPU q's
    memory and can be read AND written by this rank's kernels over NVLink.  The owner must keep `t`
    alive.  torch sub-allocates: several tensors can live in one cudaMalloc block, so a block that is
    already mapped is not opened a second time."""
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
        base = _OPEN_HANDLES.get((q, h))
        if base is None:
            ptr = ctypes.c_void_p()
            with torch.cuda.device(t.device):
                _lib.check_call(lib.dglb200_ipc_open(ctypes.c_char_p(h), ctypes.byref(ptr)))
            base = _OPEN_HANDLES[(q, h)] = int(ptr.value)
        if 0 in shape:
            peers.append(torch.empty(shape, dtype=t.dtype, device=t.device))
            continue
        pt = torch.as_tensor(_RawCuda(base + off, shape, typestr), device=t.device)
        peers.append(pt.view(view_as) if view_as is not None else pt)
    return peers


def close_peer_mappings():
    """Unmap every peer buffer this process opened with share_with_peers (dglb200_ipc_close).  Call it
    once the tensors returned by share_with_peers are no longer used (e.g. before tearing a
    partition down in a long-lived process); process exit unmaps them anyway."""
    import ctypes
    from . import _lib
    lib = _lib.get_lib()
    while _OPEN_HANDLES:
        _, base = _OPEN_HANDLES.popitem()
        _lib.check_call(lib.dglb200_ipc_close(ctypes.c_void_p(base)))


def release_peer_memory(group=None):
    """Collective: after every rank has dropped the exchange buffers it exported (part.release()), unmap
    the peers' buffers here.  An exported allocation stays physically allocated until EVERY importer has
    closed its mapping, so a process that rebuilds partitions (benchmarks, sweeps) must call this
    between them or the exporters' memory accumulates."""
    if dist.is_initialized():
        dist.barrier(group=group)
    if torch.cuda.is_available():
        torch.cuda.synchronize()
    close_peer_mappings()
    if torch.cuda.is_available():
        torch.cuda.empty_cache()
    if dist.is_initialized():
        dist.barrier(group=group)


class PeerSignals(object):
    """Flag words of one exchange channel: `slots` x world int32 in THIS rank's memory, mapped into
    every peer.  Word [slot, q] is written by rank q only (and only ever grows: it carries the epoch
    of the last exchange rank q completed for that slot) and read by this rank only."""

    def __init__(self, slots, rank, world, device, group=None):
        apply_tuning()
        self.rank, self.world, self.slots = rank, world, slots
        self.flags = torch.zeros((slots, world), dtype=torch.int32, device=device)
        self.counter = torch.zeros(max(slots, 1), dtype=torch.int32, device=device)
        self.err = torch.zeros(1, dtype=torch.int32, device=device)
        torch.cuda.synchronize(device) if self.flags.is_cuda else None
        self.peer_flags = share_with_peers(self.flags, rank, world, group)

    def local(self, slot):
        """My flag words of `slot`, one per rank (what peer_wait polls)."""
        return self.flags[slot]

    def remote(self, slot):
        """For every rank q (self included): address of MY word [slot, rank] inside q's flag array."""
        return [int(self.peer_flags[q][slot, self.rank:self.rank + 1].data_ptr()) for q in range(self.world)]

    def errors(self):
        return int(self.err.item())


class StreamBarrier(object):
    """Cross-rank barrier in STREAM order (no host stall): a 1-element all-reduce.  Every rank's
    kernels enqueued before it are complete before any rank's kernels enqueued after it start."""

    def __init__(self, device, group=None):
        self.token = torch.zeros(1, dtype=torch.float32, device=device)
        self.group = group

    def __call__(self):
        dist.all_reduce(self.token, group=self.group)


class _Channel(object):
    """State of one exchange buffer: the buffer itself, its peer mappings, its flag words and the
    epoch counter of the signal transport.  One per (key, width, dtype): a model that aggregates at
    several widths (a 2-layer GCN) never reallocates a buffer that peers have mapped."""

    def __init__(self, buf):
        self.buf = buf
        self.peer = None          # every rank's buffer mapped here
        self.wire = None          # bf16 staging of the halo rows (wire mode)
        self.peer_wire = None
        self.sig = None           # PeerSignals
        self.epoch = 0
        self.pending_free = 0     # epoch whose "buffer free" signal is still owed (exchange() without SpMM)


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
        self._spmm_is_default = spmm_fn is None
        self._splits = {}         # id(csr) -> (own-column CSR, remote-column CSR)
        self._full = {}           # (key, feat, dtype) -> _Channel
        self._last = {}           # key -> channel of the last exchange under that key
        self._barrier = None
        self._comm = None
        self._peer_on = False

    # -- buffers ---------------------------------------------------------------------------------
    def _channel(self, key, feat, like):
        ck = (key, int(feat), like.dtype)
        ch = self._full.get(ck)
        if ch is None:
            if self._peer_on and self.world > 1:
                raise DGLError("RowPartition: no peer-mapped exchange buffer for key %r at width %d (%s); call "
                               "enable_peer_transport(%d, like, keys=(%r,)) collectively first"
                               % (key, feat, like.dtype, feat, key))
            ch = self._full[ck] = _Channel(torch.empty((self.n, feat), dtype=like.dtype, device=like.device))
        self._last[key] = ch
        return ch

    def enable_peer_transport(self, feat, like, keys=("x", "dy")):
        """Exchange over NVLink peer memory instead of NCCL: every rank's gather buffer is mapped into
        every other rank (CUDA IPC).  Collective: every rank calls it with the same arguments."""
        if self.world == 1:
            return
        self._peer_on = False
        for key in keys:
            ch = self._channel(key, feat, like)
            if ch.peer is None:
                ch.peer = share_with_peers(ch.buf, self.rank, self.world, self.group)
                # the push kernel moves 16-byte units: widths whose rows are not 16-byte multiples (41
                # classes) keep the round-1 transport (block copies between stream barriers)
                if PULL_MODE == "signal" and (feat * ch.buf.element_size()) % 16 == 0:
                    n_slab = self._slabs(feat, like.dtype)
                    ch.sig = PeerSignals(len(n_slab) + 1, self.rank, self.world, like.device, self.group)
                elif self._barrier is None:
                    self._barrier = StreamBarrier(like.device, self.group)
        self._peer_on = True

    def _global_rank(self, p):
        return p if self.group is None else dist.get_global_rank(self.group, p)

    def full_buffer(self, key, feat, like):
        """Persistent [n, feat] gather buffer (one per key, width and dtype)."""
        return self._channel(key, feat, like).buf

    def views(self, full):
        return [full[self.bounds[p]:self.bounds[p + 1]] for p in range(self.world)]

    def release(self):
        """Drop the exchange buffers (the peer mappings stay open until release_peer_memory())."""
        self._full.clear()
        self._last.clear()
        self._peer_on = False

    def comm_errors(self):
        """Signal waits that timed out since the start (0 = none) + out-of-range ids seen by gathers."""
        ops.check_bad_ids()
        return sum(ch.sig.errors() for ch in self._full.values() if ch.sig is not None)

    def local_spmm(self, csr, x, out):
        """The local kernel alone over an already exchanged operand (benchmark kernel-only timing)."""
        return self._spmm(csr, x, out)

    # -- all-gather --------------------------------------------------------------------------------
    def all_gather_rows(self, local, key):
        """[rows, feat] block of every rank -> [n, feat] on every rank (uneven block sizes)."""
        ch = self._channel(key, local.shape[1], local)
        full = ch.buf
        views = self.views(full)
        if self.world == 1:
            if views[0].data_ptr() != local.data_ptr():
                views[0].copy_(local)
            return full
        if views[self.rank].data_ptr() != local.data_ptr():
            views[self.rank].copy_(local)
        if ch.sig is not None:
            self._signal_push(ch, self._slabs(local.shape[1], local.dtype), on_comm=False)
            for i in range(ch.sig.slots - 1):
                ops.peer_wait(ch.sig.local(i), ch.epoch, ch.sig.err)
            ch.pending_free = ch.epoch        # the caller's SpMM follows on this stream
            return full
        if ch.peer is not None:
            self._barrier()                                   # every owner's block is written
            for k in range(1, self.world):                    # start at a different owner on every rank
                q = (self.rank + k) % self.world
                if views[q].numel():
                    views[q].copy_(ch.peer[q][self.bounds[q]:self.bounds[q + 1]], non_blocking=True)
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

    SLAB = 0      # round-1 feature-slab PULL pipeline (transport "copy"): 0 = off, measured slower

    def _slabs(self, feat, dtype):
        """Column blocks [(s0, width)] of the pipelined push all-gather: the SpMM's own L2 slab width
        (64 fp32 at C2), so that the per-slab launches do the work of the one slabbed launch."""
        es = torch.empty(0, dtype=dtype).element_size()
        if ALLGATHER_SLABS:               # explicit, possibly unequal widths (a narrow first slab lands sooner)
            ws = [int(v) for v in ALLGATHER_SLABS.split(",")]
            if sum(ws) == feat and all(v > 0 and (v * es) % 16 == 0 for v in ws):
                out, s0 = [], 0
                for v in ws:
                    out.append((s0, v))
                    s0 += v
                return out
        w = ALLGATHER_SLAB if ALLGATHER_SLAB > 0 else (64 if self.world <= 2 else 128)
        if w <= 0 or feat <= w or feat % w or (w * es) % 16:
            return [(0, feat)]
        return [(s0, w) for s0 in range(0, feat, w)]

    def _split(self, csr):
        """The row block cut by SOURCE owner: (edges whose source this rank owns, all other edges), both
        over the global column space and with every row's slot order kept.  Plan-time integer work."""
        key = id(csr)
        sp = self._splits.get(key)
        if sp is None:
            cols = csr.indices
            own = (cols >= self.r0) & (cols < self.r1)
            rows = ops.csr_row_ids(csr)
            parts = []
            for m in (own, ~own):
                cnt = torch.bincount(rows[m], minlength=csr.n_rows)
                indptr = torch.zeros(csr.n_rows + 1, dtype=torch.int64, device=cols.device)
                indptr[1:] = torch.cumsum(cnt, 0)
                sub = ops.CSR(csr.n_rows, csr.n_cols, indptr, cols[m].contiguous(),
                              csr.eids[m].contiguous() if csr.eids is not None else None)
                sub.plan()
                parts.append(sub)
            sp = self._splits[key] = (parts[0], parts[1], csr)     # keep csr alive: id() is the key
        return sp

    def _signal_push(self, ch, slabs, on_comm=True):
        """Enqueue the pushes of a new epoch: for every column slab ONE launch stores this rank's block
        into every peer's gather buffer and sets the slab's flag word on every rank.  The first launch
        waits (in the kernel) until every rank has released the buffer of the previous epoch."""
        if ch.pending_free:       # a bare exchange() left the release of its epoch to the next call
            ops.peer_signal(ch.sig.remote(ch.sig.slots - 1), ch.pending_free, ch.buf.device)
            ch.pending_free = 0
        ch.epoch += 1
        own = ch.buf[self.r0:self.r1]
        es = own.element_size()
        counts = [self.rows if q != self.rank else 0 for q in range(self.world)]
        free_slot = ch.sig.slots - 1
        cur = torch.cuda.current_stream(ch.buf.device)
        if on_comm:
            if self._comm is None:
                self._comm = torch.cuda.Stream(device=ch.buf.device, priority=COMM_PRIORITY)
            self._comm.wait_stream(cur)
        ctx = torch.cuda.stream(self._comm) if on_comm else _nullcontext()
        with ctx:
            for i, (s0, w) in enumerate(slabs):
                dst = [int(ch.peer[q][self.r0:self.r1, s0:s0 + w].data_ptr()) if (q != self.rank and self.rows) else 0
                       for q in range(self.world)]
                ops.peer_push(own[:, s0:s0 + w], None, counts, dst, ch.buf.stride(0) * es, w * es,
                              signal_addrs=ch.sig.remote(i), signal_value=ch.epoch, counter=ch.sig.counter[i:i + 1],
                              wait_flags=ch.sig.local(free_slot) if i == 0 else None, wait_value=ch.epoch - 1,
                              err=ch.sig.err, rotate=self.rank, max_ctas=PUSH_CTAS)

    def _aggregate(self, key, local, csr, out):
        """Exchange + SpMM.  Signal transport: the feature dimension is pipelined in slabs - slab s+1
        is pushed into the peers (NVLink stores from a side stream) while the SpMM runs over slab s
        (dglb200_spmm_csr_strided on column blocks of the gather buffer)."""
        feat = local.shape[1]
        ch = self._channel(key, feat, local)
        if ch.sig is not None:
            full = ch.buf
            own = full[self.r0:self.r1]
            if own.data_ptr() != local.data_ptr():
                own.copy_(local)
            if out is None:
                out = torch.empty((csr.n_rows, feat), dtype=local.dtype, device=local.device)
            slabs = self._slabs(feat, local.dtype)
            self._signal_push(ch, slabs)
            split = self._split(csr) if (SPLIT_LOCAL and self._spmm_is_default) else None
            if split is not None:
                # own-source edges first: their operand rows are already here, the pushes run meanwhile
                ops.spmm(split[0], ops.MSG_COPY_SRC, ops.RED_SUM, x=full, out=out)
                csr = split[1]
            for i, (s0, w) in enumerate(slabs):
                ops.peer_wait(ch.sig.local(i), ch.epoch, ch.sig.err)
                xs, os_ = (full, out) if len(slabs) == 1 else (full[:, s0:s0 + w], out[:, s0:s0 + w])
                if split is not None:
                    ops.spmm(csr, ops.MSG_COPY_SRC, ops.RED_SUM, x=xs, out=os_, accumulate=True)
                else:
                    self._spmm(csr, xs, os_)
            ops.peer_signal(ch.sig.remote(ch.sig.slots - 1), ch.epoch, full.device)   # buffer released
            torch.cuda.current_stream(full.device).wait_stream(self._comm)   # my block may be rewritten only after my pushes
            return out
        if ch.peer is None or self.SLAB <= 0 or feat <= self.SLAB or feat % self.SLAB:
            return self._spmm(csr, self.exchange(key, local), out)
        full = ch.buf
        views = self.views(full)
        if views[self.rank].data_ptr() != local.data_ptr():
            views[self.rank].copy_(local)
        if out is None:
            out = torch.empty((csr.n_rows, feat), dtype=local.dtype, device=local.device)
        if self._comm is None:
            self._comm = torch.cuda.Stream(device=local.device, priority=COMM_PRIORITY)
        cur = torch.cuda.current_stream(local.device)
        peers = ch.peer
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
        """y_local = (A x)[r0:r1] for update_all(copy_src, sum).  `key` names the gather buffer."""
        return self._aggregate(key, x_local, self.loc_in, out)

    def backward(self, dy_local, out=None, key="dy"):
        """dx_local = (A^T dy)[r0:r1]."""
        return self._aggregate(key, dy_local, self.loc_out, out)


class _nullcontext(object):
    def __enter__(self):
        return None

    def __exit__(self, *a):
        return False


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
    Per call (signal transport), for k = 0..chunks-1 on a side stream: ONE launch pushes the rows of
    chunk k that every peer needs into the peers' compact buffers and sets their flag words; the
    SpMM over chunk k's rows waits on this rank's own flag words, so it starts as soon as ITS halo has
    landed, while chunk k+1 is in flight.
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
        fr = chunk_fractions(self.chunks)
        tgt = torch.tensor(fr[:-1], dtype=torch.float64, device=dev) * float(ip[-1])
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
            # the tables the gather / push kernels index with are validated ONCE here (the kernels skip
            # an out-of-range id and raise a flag that nothing on the hot path reads)
            if serve.numel() and (int(serve.min()) < 0 or int(serve.max()) >= max(self.rows, 1)):
                raise DGLError("HaloExchange: a peer asked rank %d for a row outside its block [0, %d)"
                               % (rank, self.rows))
            self.send_counts.append(sc)
            self.need_cat.append(need_cat.contiguous())
            self.serve_local.append(serve.contiguous())           # already local ids of THIS rank
        # push transports: where the rows I serve land inside requester r's compact buffer, per chunk
        self.push_dst = None
        self._rows_of_rank = [self.rows]
        if world > 1:
            objs = [None] * world
            dist.all_gather_object(objs, (self.chunk_start, self.recv_counts, self.rows), group=group)
            self.push_dst = [[int(objs[r][0][k] + sum(objs[r][1][k][:rank])) for r in range(world)]
                             for k in range(self.chunks)]
            self._rows_of_rank = [int(objs[r][2]) for r in range(world)]
        self.chunk_csr = None
        self.timing = []
        self.timeline = []        # TIMING: per op [op start, wait_0 done, spmm_0 done, wait_1 done, ...] events (compute stream)
        self._chan = {}           # (feat, dtype) -> _Channel
        self._cur = None          # channel of the last call
        self._peer_on = False
        self._barrier = None
        self._comm = None

    # -- buffers / transports -----------------------------------------------------------------
    def _channel(self, feat, like):
        ck = (int(feat), like.dtype)
        ch = self._chan.get(ck)
        if ch is None:
            if self._peer_on and self.world > 1:
                raise DGLError("HaloExchange: no peer-mapped buffer at width %d (%s); call "
                               "enable_peer_transport(%d, like) collectively first" % (feat, like.dtype, feat))
            ch = self._chan[ck] = _Channel(torch.empty((self.rows + self.n_halo, feat), dtype=like.dtype,
                                                       device=like.device))
        self._cur = ch
        return ch

    def buffer(self, feat, like):
        """Compact feature buffer [rows + n_halo, feat]; rows [0, rows) are this rank's block.  One
        buffer (with its peer mappings) per width and dtype."""
        return self._channel(feat, like).buf

    @property
    def _wire(self):              # the bf16 staging buffer of the channel in use (tests / stats)
        return self._cur.wire if self._cur is not None else None

    def release(self):
        self._chan.clear()
        self._cur = None
        self._peer_on = False

    def errors(self):
        return sum(ch.sig.errors() for ch in self._chan.values() if ch.sig is not None)

    def enable_peer_transport(self, feat, like):
        """Move the halo rows over NVLink peer memory (no NCCL on the data path).  Collective."""
        if self.world == 1:
            return
        self._peer_on = False
        ch = self._channel(feat, like)
        if ch.peer is None:
            ch.peer = share_with_peers(ch.buf, self.rank, self.world, self.group)
            if WIRE == "bf16" and PULL_MODE in ("push", "signal") and ch.buf.dtype == torch.float32:
                # halo rows arrive here as bf16 (written by their owners) and are widened into `buf`
                ch.wire = torch.empty((max(self.n_halo, 1), feat), dtype=torch.bfloat16, device=like.device)
                ch.peer_wire = share_with_peers(ch.wire, self.rank, self.world, self.group)
            unit = 32 if ch.wire is not None else 16
            if PULL_MODE == "signal" and (feat * ch.buf.element_size()) % unit == 0:
                ch.sig = PeerSignals(self.chunks + 1, self.rank, self.world, like.device, self.group)
            elif self._barrier is None:
                self._barrier = StreamBarrier(like.device, self.group)
        self._peer_on = True

    def row_chunks(self):
        """The remapped CSR cut into the plan's row chunks (views; each with its own long-row plan)."""
        if self.chunk_csr is None:
            self.chunk_csr = [ops.csr_row_block(self.csr, self.row_cuts[k], self.row_cuts[k + 1])
                              for k in range(self.chunks)]
        return self.chunk_csr

    def halo_bytes(self, feat, elem=4):
        return self.n_halo * feat * elem

    # -- data path ---------------------------------------------------------------------------
    def _push_chunk_signal(self, ch, k):
        """Signal transport: ONE launch stores chunk k's rows into every requester's compact buffer (or
        its bf16 staging buffer) and sets the chunk's flag word on every rank."""
        buf = ch.buf
        es = buf.element_size()
        sc = self.send_counts[k]
        feat = buf.shape[1]
        dst = []
        for r in range(self.world):
            if r == self.rank or not sc[r]:
                dst.append(0)
            elif ch.peer_wire is not None:
                h0 = self.push_dst[k][r] - self._rows_of_rank[r]          # position inside r's halo region
                dst.append(int(ch.peer_wire[r][h0:h0 + 1].data_ptr()))
            else:
                d0 = self.push_dst[k][r]
                dst.append(int(ch.peer[r][d0:d0 + 1].data_ptr()))
        to_bf16 = ch.peer_wire is not None
        free_slot = ch.sig.slots - 1
        ops.peer_push(buf[:self.rows], self.serve_local[k], sc, dst,
                      feat * (2 if to_bf16 else es), feat * es,
                      signal_addrs=ch.sig.remote(k), signal_value=ch.epoch, counter=ch.sig.counter[k:k + 1],
                      wait_flags=ch.sig.local(free_slot) if k == 0 else None, wait_value=ch.epoch - 1,
                      err=ch.sig.err, to_bf16=to_bf16, rotate=self.rank, max_ctas=PUSH_CTAS)

    def _fetch_chunk(self, k, ch):
        """Bring chunk k's halo rows into the buffer (current stream; barrier / NCCL transports)."""
        buf = ch.buf
        start = self.chunk_start[k]
        cnts = self.recv_counts[k]
        if ch.peer is not None and PULL_MODE in ("push", "signal"):
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
                    if ch.peer_wire is not None:
                        h0 = d0 - self._rows_of_rank[r]          # position inside r's halo region
                        ops.gather_rows_convert(buf[:self.rows], self.serve_local[k][seg[r]:seg[r + 1]],
                                                out=ch.peer_wire[r][h0:h0 + sc[r]], max_ctas=ctas)
                    else:
                        ops.gather_rows(buf[:self.rows], self.serve_local[k][seg[r]:seg[r + 1]],
                                        out=ch.peer[r][d0:d0 + sc[r]], max_ctas=ctas)
            return
        if ch.peer is not None:
            # pull this chunk's rows out of the owners' memory (NVLink peer loads)
            seg = [0]
            for q in range(self.world):
                seg.append(seg[-1] + cnts[q])
            if seg[-1] and PULL_MODE == "perpeer":
                for j in range(1, self.world):
                    q = (self.rank + j) % self.world
                    if cnts[q]:
                        ops.gather_rows(ch.peer[q][:self.bounds[q + 1] - self.bounds[q]], self.need_local[k][q],
                                        out=buf[start + seg[q]:start + seg[q + 1]], max_ctas=PULL_CTAS)
            elif seg[-1]:
                srcs = [ch.peer[q][:self.bounds[q + 1] - self.bounds[q]] if q != self.rank else None
                        for q in range(self.world)]
                ops.gather_rows_multi(srcs, seg, self.need_cat[k], buf[start:start + seg[-1]], max_ctas=PULL_CTAS)
            return
        send = self._gather(buf[:self.rows], self.serve_local[k])
        dist.all_to_all_single(buf[start:start + sum(cnts)], send, output_split_sizes=cnts,
                               input_split_sizes=self.send_counts[k], group=self.group)

    def _widen_chunk(self, k, ch):
        """bf16 wire mode: chunk k's halo rows have landed in the bf16 staging buffer; widen them
        into the fp32 operand (local HBM traffic only)."""
        if ch.wire is None:
            return
        cnt = sum(self.recv_counts[k])
        if cnt:
            h0 = self.chunk_start[k] - self.rows
            ops.gather_rows_convert(ch.wire[h0:h0 + cnt], None,
                                    out=ch.buf[self.chunk_start[k]:self.chunk_start[k] + cnt])

    def _mixed(self, ch):
        return ch.wire is not None and MIXED_HALO and ch.buf.shape[1] % 8 == 0

    def chunk_spmm(self, ch, k, csr_k, out_k, spmm_fn):
        """The SpMM over one row chunk once its halo has landed.  bf16 wire: the kernel reads own rows
        from the fp32 block and halo rows from the bf16 staging buffer (k = None: the whole block)."""
        if self._mixed(ch):
            return ops.spmm(csr_k, ops.MSG_COPY_SRC, ops.RED_SUM, x=ch.buf[:self.rows], out=out_k,
                            x_halo=ch.wire, halo_first=self.rows)[0]
        if ch.wire is not None and k is not None:
            self._widen_chunk(k, ch)
        return spmm_fn(csr_k, ch.buf, out_k)

    def _own_rows(self, ch, x_local):
        own = ch.buf[:self.rows]
        if own.data_ptr() != x_local.data_ptr():
            own.copy_(x_local)

    def _start_epoch(self, ch):
        if ch.pending_free:       # a bare exchange() left the release of its epoch to the next call
            ops.peer_signal(ch.sig.remote(ch.sig.slots - 1), ch.pending_free, ch.buf.device)
            ch.pending_free = 0
        ch.epoch += 1

    def exchange(self, x_local):
        """Unpipelined: returns the compact buffer with own rows and every halo row."""
        ch = self._channel(x_local.shape[1], x_local)
        buf = ch.buf
        self._own_rows(ch, x_local)
        if self.world == 1:
            return buf
        if ch.sig is not None:
            self._start_epoch(ch)
            for k in range(self.chunks):
                self._push_chunk_signal(ch, k)
            for k in range(self.chunks):
                ops.peer_wait(ch.sig.local(k), ch.epoch, ch.sig.err)
                self._widen_chunk(k, ch)
            ch.pending_free = ch.epoch        # the caller's SpMM follows on this stream
            return buf
        if self._barrier is not None:
            self._barrier()
        if TIMING and x_local.is_cuda:
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
        for k in range(self.chunks):
            self._fetch_chunk(k, ch)
        if TIMING and x_local.is_cuda:
            b.record()
            self.timing.append((a, b))
        if self._barrier is not None:
            self._barrier()
        for k in range(self.chunks):
            self._widen_chunk(k, ch)
        return buf

    def aggregate(self, x_local, spmm_fn, out):
        """Pipelined exchange + SpMM: y[rows of chunk k] is computed while chunk k+1's halo is in
        flight on the side stream.  `out` [rows, feat] must be given."""
        ch = self._channel(x_local.shape[1], x_local)
        if self.world > 1 and ch.sig is not None:
            return self._aggregate_signal(ch, x_local, spmm_fn, out)
        if self.world == 1 or self.chunks == 1:
            return spmm_fn(self.csr, self.exchange(x_local), out)
        if ch.peer is not None and PULL_MODE in ("push", "signal"):
            return self._aggregate_pushed(ch, x_local, spmm_fn, out)
        buf = ch.buf
        self._own_rows(ch, x_local)
        if not x_local.is_cuda and ch.peer is None:               # CPU (gloo tests): same schedule, no streams
            for k, csr_k in enumerate(self.row_chunks()):
                self._fetch_chunk(k, ch)
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
                self._fetch_chunk(k, ch)
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

    def _aggregate_signal(self, ch, x_local, spmm_fn, out):
        """Signal transport.  Side stream: push chunk 0 (waits in the kernel until every rank has
        released the buffer of the previous epoch), push chunk 1, ...  Compute stream: wait for chunk
        k's flag words, [widen], SpMM over chunk k's rows, ...; finally tell every rank that this
        rank's buffer is free again.  A pushed chunk is complete on the receiver when EVERY owner's flag
        word carries the epoch."""
        buf = ch.buf
        self._own_rows(ch, x_local)
        if self._comm is None:
            self._comm = torch.cuda.Stream(device=buf.device, priority=COMM_PRIORITY)
        cur = torch.cuda.current_stream(buf.device)
        self._start_epoch(ch)
        self._comm.wait_stream(cur)                               # my rows are written
        with torch.cuda.stream(self._comm):
            for k in range(self.chunks):
                if TIMING:
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record()
                self._push_chunk_signal(ch, k)
                if TIMING:
                    b.record()
                    self.timing.append((a, b))
        chunk_csrs = self.row_chunks() if self.chunks > 1 else [self.csr]
        tl = [] if TIMING else None
        if TIMING:
            e0 = torch.cuda.Event(enable_timing=True)
            e0.record()
            tl.append(e0)
        for k, csr_k in enumerate(chunk_csrs):
            ops.peer_wait(ch.sig.local(k), ch.epoch, ch.sig.err)
            if TIMING:
                ew = torch.cuda.Event(enable_timing=True)
                ew.record()
                tl.append(ew)
            if csr_k.n_rows:
                self.chunk_spmm(ch, k, csr_k, out[self.row_cuts[k]:self.row_cuts[k + 1]] if self.chunks > 1 else out,
                                spmm_fn)
            elif ch.wire is not None and not self._mixed(ch):
                self._widen_chunk(k, ch)
            if TIMING:
                es = torch.cuda.Event(enable_timing=True)
                es.record()
                tl.append(es)
        if TIMING:
            self.timeline.append(tl)
        ops.peer_signal(ch.sig.remote(ch.sig.slots - 1), ch.epoch, buf.device)    # my buffer is free again
        cur.wait_stream(self._comm)        # my block may be rewritten only after my pushes have read it
        return out

    def _aggregate_pushed(self, ch, x_local, spmm_fn, out):
        """Round-1 pipelined PUSH exchange + SpMM (per-peer launches, NCCL stream barriers per chunk)."""
        buf = ch.buf
        self._own_rows(ch, x_local)
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
                self._fetch_chunk(k, ch)
                if TIMING:
                    b.record()
                    self.timing.append((a, b))
                self._barrier()               # chunk k's halo has landed everywhere
                self._widen_chunk(k, ch)
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

    def local_spmm(self, csr, x, out):
        h = self.halo_in if csr is self.loc_in else self.halo_out
        if h._cur is not None and h._mixed(h._cur):
            return h.chunk_spmm(h._cur, None, csr, out, self._spmm)
        return self._spmm(csr, x, out)

    def release(self):
        self.halo_in.release()
        self.halo_out.release()
        super(HaloPartition, self).release()

    def comm_errors(self):
        ops.check_bad_ids()
        return self.halo_in.errors() + self.halo_out.errors()

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
