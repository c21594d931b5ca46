"""-m "not gpu": the PEER transports of the multi-GPU exchange (dgl_b200/dist.py) - NVLink pulls in
their three modes, the pipelined PUSH that is the default for the compacted halo, the bf16 wire mode,
and the peer all-gather of RowPartition - run with the ranks of a partition as host THREADS.

On the GPUs every rank maps the others' exchange buffers with CUDA IPC and the gather kernels read or
write peer memory directly.  Here the "peers" are threads of one process, so a peer's buffer is simply
its tensor; torch.distributed is replaced by a thread rendezvous with the same call signatures; the
kernels are the emulated CUDA sources of tests/native/cuda_emu (see tests/test_kernels_host_emu.py).
What is checked is everything the gloo tests of tests/test_partition.py cannot reach: the push offsets
(`push_dst`), the per-chunk barriers, the rotating peer order, the wire staging buffer and its widening,
against the single-rank SpMM - bit for bit with the fp32 wire (the remapped CSR keeps every row's slot
order), within the bf16 bar (1e-2) with the bf16 wire.
"""
import contextlib
import threading

import pytest
import torch

from tests.test_kernels_host_emu import emu, emu_lib  # noqa: F401  (fixtures)
from tests.util import random_graph

pytestmark = pytest.mark.timeout(600)


class ThreadDist(object):
    """Stand-in for torch.distributed inside dgl_b200.dist: ranks are threads of this process."""

    class ReduceOp(object):
        SUM, MAX, MIN = "sum", "max", "min"

    def __init__(self, world):
        self.world = world
        self._bar = threading.Barrier(world)
        self._slots = [None] * world
        self._tl = threading.local()

    def set_rank(self, r):
        self._tl.rank = r

    def is_initialized(self):
        return True

    def get_world_size(self, group=None):
        return self.world

    def get_rank(self, group=None):
        return self._tl.rank

    def get_backend(self, group=None):
        return "threads"

    def get_global_rank(self, group, p):
        return p

    def _swap(self, obj):
        """Every rank publishes obj; returns the list of all ranks' objects."""
        self._slots[self._tl.rank] = obj
        self._bar.wait()
        got = list(self._slots)
        self._bar.wait()
        return got

    def all_gather_object(self, objs, obj, group=None):
        objs[:] = self._swap(obj)

    def all_reduce(self, t, op="sum", group=None):
        vals = self._swap(t.clone())
        acc = vals[0].clone()
        for v in vals[1:]:
            acc = acc + v if op == "sum" else (torch.maximum(acc, v) if op == "max" else torch.minimum(acc, v))
        t.copy_(acc)

    def all_to_all_single(self, out, inp, output_split_sizes=None, input_split_sizes=None, group=None):
        w = self.world
        ins = input_split_sizes if input_split_sizes is not None else [inp.shape[0] // w] * w
        pub = self._swap((inp, list(ins)))
        me, parts = self._tl.rank, []
        for q in range(w):
            t, sizes = pub[q]
            o = sum(sizes[:me])
            parts.append(t[o:o + sizes[me]])
        res = torch.cat(parts) if parts else inp[:0]
        out[:res.shape[0]].copy_(res)
        self._bar.wait()            # nobody reuses its input before everyone has read it

    def all_gather(self, outs, t, group=None):
        for o, v in zip(outs, self._swap(t)):
            o.copy_(v)

    def broadcast(self, t, src=0, group=None, async_op=False):
        t.copy_(self._swap(t)[src])

        class _Done(object):
            def wait(self):
                return None
        return _Done()


class _Stream(object):
    def wait_stream(self, other):
        pass

    def wait_event(self, ev):
        pass


class _Event(object):
    def __init__(self, *a, **k):
        pass

    def record(self, *a, **k):
        pass


@pytest.fixture()
def peers(emu, monkeypatch):
    """dgl_b200.dist wired for threads-as-ranks on the emulated kernels."""
    from dgl_b200 import dist as dmod
    monkeypatch.setattr(torch.cuda, "Stream", lambda *a, **k: _Stream())
    monkeypatch.setattr(torch.cuda, "current_stream", lambda *a, **k: _Stream())
    monkeypatch.setattr(torch.cuda, "stream", lambda s: contextlib.nullcontext())
    monkeypatch.setattr(torch.cuda, "Event", _Event)

    def make(world):
        td = ThreadDist(world)
        monkeypatch.setattr(dmod, "dist", td)
        # a peer's buffer is the peer thread's tensor itself
        monkeypatch.setattr(dmod, "share_with_peers",
                            lambda t, rank, world_, group=None: td._swap(t))
        return td
    return make


def _graph(n, e, seed):
    from dgl_b200 import ops
    src, dst = random_graph(n, e, seed, hubs=2, hub_deg=n // 3, zero_deg_frac=0.15)
    s, d = torch.from_numpy(src), torch.from_numpy(dst)
    return ops.csr_build(d, s, n, n), ops.csr_build(s, d, n, n)


def _reference(csr_in, csr_out, x, dy):
    from dgl_b200 import ops
    y = ops.spmm(csr_in, ops.MSG_COPY_SRC, ops.RED_SUM, x=x)[0]
    dx = ops.spmm(csr_out, ops.MSG_COPY_SRC, ops.RED_SUM, x=dy)[0]
    return y, dx


def _run(world, td, per_rank):
    errs = [None] * world

    def body(r):
        td.set_rank(r)
        try:
            per_rank(r)
        except BaseException as e:  # noqa: BLE001
            errs[r] = e
            td._bar.abort()
    ts = [threading.Thread(target=body, args=(r,)) for r in range(world)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    real = [e for e in errs if e is not None and not isinstance(e, threading.BrokenBarrierError)]
    if real:
        raise real[0]
    if any(errs):
        raise [e for e in errs if e is not None][0]


@pytest.mark.parametrize("mode,chunks,wire", [("signal", 3, "f32"), ("signal", 1, "f32"), ("signal", 2, "bf16"),
                                              ("push", 2, "f32"), ("push", 1, "f32"), ("perpeer", 1, "f32"),
                                              ("fused", 2, "f32"), ("push", 3, "bf16")])
def test_halo_partition_peer_transports(peers, monkeypatch, mode, chunks, wire):
    from dgl_b200 import dist as dmod
    world, n, feat = 3, 90, 16
    monkeypatch.setattr(dmod, "PULL_MODE", mode)
    monkeypatch.setattr(dmod, "WIRE", wire)
    csr_in, csr_out = _graph(n, 700, 5)
    gen = torch.Generator().manual_seed(1)
    x, dy = torch.randn(n, feat, generator=gen), torch.randn(n, feat, generator=gen)
    want_y, want_dx = _reference(csr_in, csr_out, x, dy)
    td = peers(world)
    got_y, got_dx, stats = [None] * world, [None] * world, [None] * world

    def per_rank(r):
        part = dmod.HaloPartition(csr_in, csr_out, rank=r, world=world, chunks=chunks)
        like = torch.empty(0)
        part.enable_peer_transport(feat, like)
        xl = part.local_view("x", feat, like)
        dl = part.local_view("dy", feat, like)
        for step in range(3):                        # buffers are reused across steps (epochs 1..3)
            xl.copy_(x[part.r0:part.r1] * (1 << step))
            dl.copy_(dy[part.r0:part.r1] * (1 << step))
            y = part.forward(xl)
            dx = part.backward(dl)
        got_y[r], got_dx[r] = (part.r0, y.clone()), (part.r0, dx.clone())
        stats[r] = (part.halo_in.n_halo, part.exchange_bytes(feat), part.halo_in._wire is not None)
    _run(world, td, per_rank)
    y = torch.cat([t for _, t in sorted(got_y, key=lambda p: p[0])])
    dx = torch.cat([t for _, t in sorted(got_dx, key=lambda p: p[0])])
    if wire == "f32":
        assert torch.equal(y, 4 * want_y) and torch.equal(dx, 4 * want_dx)     # same slot order: bit-exact
    else:
        assert all(s[2] for s in stats)              # the staging buffer exists on every rank
        for got, want in ((y, 4 * want_y), (dx, 4 * want_dx)):
            tol = 1e-2 * float(want.abs().max())
            assert float((got - want).abs().max()) <= tol
            assert not torch.equal(got, want)        # the halo really went through bf16
    assert sum(s[0] for s in stats) > 0              # there was a halo to exchange


@pytest.mark.parametrize("mode,feat,slab", [("signal", 8, 4), ("signal", 8, 64), ("signal", 5, 64), ("push", 8, 64),
                                            ("signal", 16, "4,12")])
def test_row_partition_peer_all_gather(peers, monkeypatch, mode, feat, slab):
    """RowPartition over peer memory: the slab-pipelined push + flag signals (4-wide slabs of an 8-wide
    row, one slab, a width the 16-byte push cannot move -> barrier transport) and round 1's block copies."""
    from dgl_b200 import dist as dmod
    monkeypatch.setattr(dmod, "PULL_MODE", mode)
    if isinstance(slab, str):                    # explicit unequal slab widths (DGL_B200_AG_SLABS)
        monkeypatch.setattr(dmod, "ALLGATHER_SLABS", slab)
    else:
        monkeypatch.setattr(dmod, "ALLGATHER_SLAB", slab)
    world, n = 3, 70
    csr_in, csr_out = _graph(n, 500, 9)
    gen = torch.Generator().manual_seed(2)
    x, dy = torch.randn(n, feat, generator=gen), torch.randn(n, feat, generator=gen)
    want_y, want_dx = _reference(csr_in, csr_out, x, dy)
    td = peers(world)
    got = [None] * world

    def per_rank(r):
        part = dmod.RowPartition(csr_in, csr_out, rank=r, world=world)
        like = torch.empty(0)
        part.enable_peer_transport(feat, like)
        xl = part.local_view("x", feat, like)
        dl = part.local_view("dy", feat, like)
        for step in range(3):
            xl.copy_(x[part.r0:part.r1] * (1 << step))
            dl.copy_(dy[part.r0:part.r1] * (1 << step))
            y, dx = part.forward(xl).clone(), part.backward(dl).clone()
        got[r] = (part.r0, y / 4, dx / 4)
        assert part.comm_errors() == 0
    _run(world, td, per_rank)
    got.sort(key=lambda p: p[0])
    for have, want in ((torch.cat([g[1] for g in got]), want_y), (torch.cat([g[2] for g in got]), want_dx)):
        # signal transport: own-source block first, remote block accumulated (one more rounding)
        assert float((have - want).abs().max()) <= 1e-5 * float(want.abs().max())
        if mode != "signal" or feat % 4:
            assert torch.equal(have, want)


def test_two_widths_keep_their_own_peer_buffers(peers, monkeypatch):
    """A 2-layer model aggregates at two widths through one partition: every (key, width) owns its
    exchange buffer and peer mapping, nothing is reallocated under a live mapping (round-1 advisor
    finding), and an unmapped width raises instead of silently exchanging through a stale mapping."""
    from dgl_b200 import dist as dmod
    from dgl_b200 import DGLError
    monkeypatch.setattr(dmod, "PULL_MODE", "signal")
    world, n = 3, 80
    csr_in, csr_out = _graph(n, 600, 11)
    gen = torch.Generator().manual_seed(3)
    xs = {8: torch.randn(n, 8, generator=gen), 4: torch.randn(n, 4, generator=gen)}
    want = {f: _reference(csr_in, csr_out, x, x) for f, x in xs.items()}
    td = peers(world)
    got = {8: [None] * world, 4: [None] * world}
    raised = [False] * world

    def per_rank(r):
        for cls in (dmod.HaloPartition, dmod.RowPartition):
            part = cls(csr_in, csr_out, rank=r, world=world)
            like = torch.empty(0)
            part.enable_peer_transport(8, like)
            part.enable_peer_transport(4, like)
            for _ in range(2):
                for f in (8, 4, 8):
                    xl = xs[f][part.r0:part.r1].contiguous()
                    y = part.forward(xl).clone()
                    dx = part.backward(xl).clone()
                    got[f][r] = (part.r0, y, dx)
            try:
                part.forward(torch.randn(part.rows, 12))
            except DGLError:
                raised[r] = True
            assert part.comm_errors() == 0
            td._bar.wait()
            for f in (8, 4):
                ys = sorted(got[f], key=lambda p: p[0])
                for col in (1, 2):
                    have, ref = torch.cat([g[col] for g in ys]), want[f][col - 1]
                    assert float((have - ref).abs().max()) <= 1e-5 * float(ref.abs().max())
                    if cls is dmod.HaloPartition:
                        assert torch.equal(have, ref)
            td._bar.wait()
    _run(world, td, per_rank)
    assert all(raised)
