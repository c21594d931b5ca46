"""-m gpu: the kernel configurations the benchmarks time, checked at the sizes where they engage.

Round-1 review: the C2 headline runs F = 256 in four L2 slabs along blockIdx.y and C5 walks 16 rows per
lane group, but no test reached either path under an assertion.  Here: the slab dispatch and
rows_per_group forced through dglb200_set_tuning on small graphs against the oracle (and bit for bit
against the default path - "same answer by a different path", the model of the reference's
tests/compute/test_specialization.py:361-515), a C5-shaped 2M-row graph where rows_per_group = 16 is the
automatic choice, the full-size C2 graph at F = 256 through bench.py's own self-check, and a 2-rank
run of every peer transport against the single-rank SpMM (skipped with fewer than 2 GPUs).
"""
import os
import sys

import numpy as np
import pytest
import torch

from oracle import c_oracle, oracle
from tests.util import random_graph

pytestmark = pytest.mark.gpu
RTOL = 1e-5


def _tune(**kw):
    from dgl_b200 import _lib
    lib = _lib.get_lib()
    for k in ("spmm_ucap", "spmm_slab_chunks", "spmm_threads", "spmm_rows_per_group"):
        _lib.check_call(lib.dglb200_set_tuning(k.encode(), int(kw.get(k, 0))))


def _sum(src, dst, n, x):
    if c_oracle.available():
        return c_oracle.spmm_sum(src, dst, n, "copy_src", x=x)
    return oracle.update_all(src, dst, n, "copy_src", "sum", x=x)


def _close(got, want, rtol=RTOL):
    got = got.float().cpu().numpy()
    np.testing.assert_allclose(got, want, rtol=rtol, atol=rtol * (np.abs(want).max() + 1e-30))


def _csr_pair(src, dst, n):
    from dgl_b200 import ops
    s, d = torch.from_numpy(src).cuda(), torch.from_numpy(dst).cuda()
    return ops.csr_build(d, s, n, n), ops.csr_build(s, d, n, n)


@pytest.mark.parametrize("slab_chunks,threads", [(16, 128), (32, 128), (64, 256), (16, 256)])
def test_feature_slab_dispatch_forced(slab_chunks, threads):
    """F = 256 processed in slabs of 16 / 32 / 64 sixteen-byte chunks along blockIdx.y (the C2 headline
    is the 16-chunk form, chosen there by the L2-residency rule): same bits as the default path, and
    within 1e-5 of the oracle.  Long rows (slices + fix-up) and empty rows included."""
    from dgl_b200 import ops
    n, feat = 3000, 256
    src, dst = random_graph(n, 150000, 11, hubs=3, hub_deg=2500, zero_deg_frac=0.2)
    x = np.random.RandomState(1).randn(n, feat).astype(np.float32)
    cin, cout = _csr_pair(src, dst, n)
    xt = torch.from_numpy(x).cuda()
    want = _sum(src, dst, n, x)
    want_t = _sum(dst, src, n, x)
    try:
        _tune()
        y0 = ops.spmm(cin, ops.MSG_COPY_SRC, ops.RED_SUM, x=xt)[0]
        _tune(spmm_slab_chunks=slab_chunks, spmm_threads=threads, spmm_ucap=4)
        y1 = ops.spmm(cin, ops.MSG_COPY_SRC, ops.RED_SUM, x=xt)[0]
        g1 = ops.spmm_backward_x(cout, xt)
    finally:
        _tune()
    assert torch.equal(y0, y1)
    _close(y1, want)
    _close(g1, want_t)


@pytest.mark.parametrize("feat", [128, 64, 16])
@pytest.mark.parametrize("rpg", [2, 4, 8, 16])
def test_rows_per_group_forced(feat, rpg):
    """A lane group walks `rpg` consecutive row steps (the C5 kernel change): forced on a sparse graph
    with 75 % empty rows and a few long rows; same bits as one row per group, oracle within 1e-5."""
    from dgl_b200 import ops
    n = 20000
    src, dst = random_graph(n, 60000, 100 + feat + rpg, hubs=2, hub_deg=1500, zero_deg_frac=0.75)
    x = np.random.RandomState(rpg).randn(n, feat).astype(np.float32)
    cin, _ = _csr_pair(src, dst, n)
    xt = torch.from_numpy(x).cuda()
    try:
        _tune(spmm_rows_per_group=1)
        y0 = ops.spmm(cin, ops.MSG_COPY_SRC, ops.RED_SUM, x=xt)[0]
        _tune(spmm_rows_per_group=rpg)
        y1 = ops.spmm(cin, ops.MSG_COPY_SRC, ops.RED_SUM, x=xt)[0]
        m1 = ops.spmm(cin, ops.MSG_COPY_SRC, ops.RED_MAX, x=xt, want_arg=True)
        _tune(spmm_rows_per_group=1)
        m0 = ops.spmm(cin, ops.MSG_COPY_SRC, ops.RED_MAX, x=xt, want_arg=True)
    finally:
        _tune()
    assert torch.equal(y0, y1)
    assert torch.equal(m0[0], m1[0]) and torch.equal(m0[1], m1[1])
    _close(y1, _sum(src, dst, n, x))


def test_c5_shaped_graph_takes_16_rows_per_group():
    """C5's shape at 1/25 scale: RMAT, 2M nodes, 6M edges (mean degree 3, most rows empty), F = 128.
    rows_per_group = 16 is the AUTOMATIC choice here (>= 4,736 * 16 row warps); forward and gradient
    against the C oracle, plus bit-equality with one row per group."""
    from benchmarks import graphs
    from dgl_b200 import ops
    n, e, feat = 2_000_000, 6_000_000, 128
    src, dst = graphs.rmat(21, e, 4, "cuda", n=n)
    cin = ops.csr_build(dst, src, n, n)
    cout = ops.csr_build(src, dst, n, n)
    deg = ops.degrees(cin)
    assert float((deg == 0).float().mean()) > 0.5          # C5-like: most rows are empty
    gen = torch.Generator(device="cuda")
    gen.manual_seed(0)
    x = torch.randn(n, feat, device="cuda", generator=gen)
    y = ops.spmm(cin, ops.MSG_COPY_SRC, ops.RED_SUM, x=x)[0]
    dx = ops.spmm_backward_x(cout, x)
    try:
        _tune(spmm_rows_per_group=1)
        y1 = ops.spmm(cin, ops.MSG_COPY_SRC, ops.RED_SUM, x=x)[0]
    finally:
        _tune()
    assert torch.equal(y, y1)
    s, d, xh = src.cpu().numpy(), dst.cpu().numpy(), x.cpu().numpy()
    _close(y, _sum(s, d, n, xh))
    _close(dx, _sum(d, s, n, xh))


def test_c2_full_size_f256_is_what_the_bench_times():
    """The headline configuration itself (N = 232,965, E = 114.6M, F = 256, four L2 slabs): bench.py's
    self-check - slab path == column-block path bit for bit, 2 x 2,000 sampled rows (hubs included)
    against the C oracle and a float64 computation - plus A 1 == in-degrees and the adjoint identity."""
    import bench
    from dgl_b200 import ops
    import dgl_b200
    cfg = bench.workload_cfg("c2")
    n, feat = cfg["n"], cfg["feat"]
    assert feat == 256
    dev = torch.device("cuda")
    g = dgl_b200.DGLGraph()
    g.add_nodes(n)
    src, dst = bench.make_edges(cfg, dev)
    g.add_edges(src, dst)
    del src, dst
    cin, cout = g.in_csr(), g.out_csr()
    g.drop_coo()
    gen = torch.Generator(device=dev)
    gen.manual_seed(0)
    x = torch.randn(n, feat, device=dev, generator=gen)
    dy = torch.randn(n, feat, device=dev, generator=gen)
    y = ops.spmm(cin, ops.MSG_COPY_SRC, ops.RED_SUM, x=x)[0]
    dx = ops.spmm(cout, ops.MSG_COPY_SRC, ops.RED_SUM, x=dy)[0]
    res = bench.check_single_gpu(cin, cout, x, dy, y, dx, feat, with_oracle=True)
    assert res["ok"], res
    assert res["slab_path_equals_column_blocks_bitwise"] is True
    assert res.get("sampled_rows_c_oracle", True) is True
    ones = torch.ones(n, feat, device=dev)
    y1 = ops.spmm(cin, ops.MSG_COPY_SRC, ops.RED_SUM, x=ones)[0]
    deg = ops.degrees(cin).float()
    assert bool((y1 == deg[:, None]).all())
    lhs = float((y.double() * dy.double()).sum())
    rhs = float((x.double() * dx.double()).sum())
    assert abs(lhs - rhs) <= 1e-6 * max(abs(lhs), abs(rhs), 1.0)


# ---- 2 ranks on 2 GPUs: every peer transport against the single-rank SpMM -------------------------

def _two_rank_worker(rank, world, port, ret):
    try:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                          CUDA_DEVICE_MAX_CONNECTIONS="32")
        import torch.distributed as dist
        torch.cuda.set_device(rank)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
        from dgl_b200 import dist as dmod, ops
        dev = torch.device("cuda", rank)
        n, feat = 20000, 128
        src, dst = random_graph(n, 300000, 21, hubs=3, hub_deg=4000, zero_deg_frac=0.3)
        s, d = torch.from_numpy(src).to(dev), torch.from_numpy(dst).to(dev)
        cin, cout = ops.csr_build(d, s, n, n), ops.csr_build(s, d, n, n)
        gen = torch.Generator(device=dev)
        gen.manual_seed(5)                                    # same data on both ranks
        x = torch.randn(n, feat, device=dev, generator=gen)
        dy = torch.randn(n, feat, device=dev, generator=gen)
        want_y = ops.spmm(cin, ops.MSG_COPY_SRC, ops.RED_SUM, x=x)[0]
        want_dx = ops.spmm(cout, ops.MSG_COPY_SRC, ops.RED_SUM, x=dy)[0]
        like = torch.empty(0, device=dev)
        cases = [("halo", "signal", 1, "f32"), ("halo", "signal", 4, "f32"), ("halo", "signal", 3, "bf16"),
                 ("halo", "push", 4, "f32"), ("halo", "push", 2, "bf16"), ("halo", "perpeer", 1, "f32"),
                 ("halo", "fused", 2, "f32"), ("row", "signal", 0, "f32"), ("row", "copy", 0, "f32"),
                 ("halo", "nccl", 2, "f32"), ("row", "nccl", 0, "f32")]
        failures = []
        for kind, mode, chunks, wire in cases:
            dmod.PULL_MODE = mode if mode != "nccl" else "signal"
            dmod.WIRE = wire
            if kind == "halo":
                part = dmod.HaloPartition(cin, cout, rank=rank, world=world, chunks=chunks)
            else:
                part = dmod.RowPartition(cin, cout, rank=rank, world=world)
            if mode != "nccl":
                part.enable_peer_transport(feat, like)
            xl, dl = part.local_view("x", feat, like), part.local_view("dy", feat, like)
            for step in range(3):                             # buffers and flag epochs are reused
                xl.copy_(x[part.r0:part.r1] * (1 << step))
                dl.copy_(dy[part.r0:part.r1] * (1 << step))
                y = part.forward(xl)
                dx = part.backward(dl)
            torch.cuda.synchronize()
            wy, wdx = 4 * want_y[part.r0:part.r1], 4 * want_dx[part.r0:part.r1]
            if wire == "f32" and not (kind == "row" and mode == "signal"):
                # halo: the remapped CSR keeps every row's slot order -> bit-exact; row partition too
                ok = torch.equal(y, wy) and torch.equal(dx, wdx)
            elif wire == "f32":
                # own-source block first, remote block accumulated: one more rounding per element
                tol = 1e-5 * float(wy.abs().max())
                ok = float((y - wy).abs().max()) <= tol and float((dx - wdx).abs().max()) <= tol
            else:
                tol = 1e-2 * float(wy.abs().max())
                ok = float((y - wy).abs().max()) <= tol and float((dx - wdx).abs().max()) <= tol
                ok = ok and not torch.equal(y, wy)            # the halo really went through bf16
            ok = ok and part.comm_errors() == 0
            if not ok:
                failures.append((kind, mode, chunks, wire))
            part.release()
            del part, xl, dl, y, dx
            dist.barrier()
        ret[rank] = failures
        dist.destroy_process_group()
    except BaseException as err:  # noqa: BLE001
        import traceback
        ret[rank] = ["EXC: %s\n%s" % (err, traceback.format_exc())]
        raise


@pytest.mark.timeout(600)
def test_two_ranks_every_peer_transport():
    """CUDA-IPC peer memory, push + flag signals (1 / 4 chunks, bf16 wire), round 1's barrier-separated
    push / pulls, the slab-pipelined all-gather and the NCCL transports, each over 3 steps (buffer and
    epoch reuse), against the single-rank SpMM: bit for bit with the fp32 wire."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    mgr = ctx.Manager()
    ret = mgr.dict()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_two_rank_worker, args=(r, 2, port, ret)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=540)
    for p in procs:
        if p.is_alive():
            p.kill()
            pytest.fail("a rank hung")
    assert all(p.exitcode == 0 for p in procs), dict(ret)
    assert dict(ret) == {0: [], 1: []}, dict(ret)
