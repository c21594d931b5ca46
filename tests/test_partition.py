"""-m "not gpu": the N>1 path (dgl_b200/dist.py) on CPU with gloo, world_size 2 and 3: edge-balanced
row partition, uneven all-gather of feature blocks, local row-block aggregation, and that the
concatenated per-rank results equal the single-process oracle.  The local compute is injected
(oracle SpMM over the rank's CSR block) because kernels need a GPU; everything else is the code
bench.py --gpus N runs."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import oracle
from tests.util import random_graph


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _csr_cpu(src, dst, n, transpose):
    from dgl_b200 import ops
    ip, ix, ei = oracle.get_adj_csr(src, dst, n, transpose)
    return ops.CSR(n, n, torch.from_numpy(ip), torch.from_numpy(ix.astype(np.int32)),
                   torch.from_numpy(ei.astype(np.int32)))


def _oracle_block_spmm(csr, x, out):
    """Checker stand-in for the kernel: y[r] = sum over the block's row r of x[indices]."""
    rows = np.repeat(np.arange(csr.n_rows), np.diff(csr.indptr.numpy()))
    y = oracle.reduce_sum(rows, x.numpy()[csr.indices.numpy()], csr.n_rows)
    return torch.from_numpy(y)


def _worker(rank, world, port, n, e, feat, q, halo=False, chunks=1):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from dgl_b200.dist import HaloPartition, RowPartition
        src, dst = random_graph(n, e, 7, hubs=3, hub_deg=n // 2, zero_deg_frac=0.1)
        if halo:
            part = HaloPartition(_csr_cpu(src, dst, n, False), _csr_cpu(src, dst, n, True),
                                 spmm_fn=_oracle_block_spmm, gather_fn=lambda x, idx: x[idx], chunks=chunks)
            assert part.halo_in.n_halo <= n - part.rows
            for h in (part.halo_in, part.halo_out):          # push plan: my rows' slot in every requester
                exp = [[h.chunk_start[k] + sum(h.recv_counts[k][:q]) for q in range(world)]
                       for k in range(h.chunks)]
                allexp = [None] * world
                dist.all_gather_object(allexp, exp)
                for k in range(h.chunks):
                    assert h.push_dst[k] == [allexp[r][k][rank] for r in range(world)]
                    assert sum(h.send_counts[k]) == h.serve_local[k].numel()
        else:
            part = RowPartition(_csr_cpu(src, dst, n, False), _csr_cpu(src, dst, n, True),
                                spmm_fn=_oracle_block_spmm)
        rng = np.random.RandomState(3)
        x = rng.randn(n, feat).astype(np.float32)
        dy = rng.randn(n, feat).astype(np.float32)
        y_loc = part.forward(torch.from_numpy(x[part.r0:part.r1].copy()))
        dx_loc = part.backward(torch.from_numpy(dy[part.r0:part.r1].copy()))
        q.put((rank, part.bounds, part.loc_in.nnz + part.loc_out.nnz, y_loc.numpy(), dx_loc.numpy()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,halo,chunks", [(2, False, 1), (3, False, 1), (2, True, 1), (3, True, 1),
                                               (2, True, 3), (3, True, 4)])
def test_row_partition_gloo(world, halo, chunks):
    n, e, feat = 300, 5000, 12
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, e, feat, q, halo, chunks)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in range(world)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    src, dst = random_graph(n, e, 7, hubs=3, hub_deg=n // 2, zero_deg_frac=0.1)
    rng = np.random.RandomState(3)
    x = rng.randn(n, feat).astype(np.float32)
    dy = rng.randn(n, feat).astype(np.float32)
    want_y = oracle.update_all(src, dst, n, "copy_src", "sum", x=x)
    want_dx = oracle.spmm_grad(src, dst, n, dy)
    bounds = res[0][1]
    assert bounds[0] == 0 and bounds[-1] == n and all(b0 <= b1 for b0, b1 in zip(bounds, bounds[1:]))
    assert all(r[1] == bounds for r in res)
    y = np.concatenate([r[3] for r in res])
    dx = np.concatenate([r[4] for r in res])
    np.testing.assert_allclose(y, want_y, rtol=1e-5, atol=1e-5 * np.abs(want_y).max())
    np.testing.assert_allclose(dx, want_dx, rtol=1e-5, atol=1e-5 * np.abs(want_dx).max())
    loads = [r[2] for r in res]
    assert sum(loads) == 2 * len(src)
    # edge balance: no rank holds more than its share plus the largest row (a hub cannot be split
    # across ranks; inside a rank the long-row plan splits it)
    deg = np.bincount(dst, minlength=n) + np.bincount(src, minlength=n)
    assert max(loads) <= 2 * len(src) / world + deg.max()


def test_bounds_degenerate_cases():
    from dgl_b200.dist import edge_balanced_bounds
    z = torch.zeros(11, dtype=torch.int64)
    b = edge_balanced_bounds(z, z, 4)          # no edges: balanced by rows
    assert b[0] == 0 and b[-1] == 10 and b == sorted(b) and max(b1 - b0 for b0, b1 in zip(b, b[1:])) <= 3
    ip = torch.tensor([0, 100, 100, 100, 100], dtype=torch.int64)      # one hub row
    b = edge_balanced_bounds(ip, ip, 2)
    assert b[0] == 0 and b[-1] == 4 and b == sorted(b)


def _worker_autograd(rank, world, port, n, e, feat, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from dgl_b200 import dist as b200dist
        src, dst = random_graph(n, e, 9, hubs=2, hub_deg=n // 3, zero_deg_frac=0.1)
        part = b200dist.RowPartition(_csr_cpu(src, dst, n, False), _csr_cpu(src, dst, n, True),
                                     spmm_fn=_oracle_block_spmm)
        rng = np.random.RandomState(5)
        x = rng.randn(n, feat).astype(np.float32)
        c = rng.randn(n, 5).astype(np.float32)
        xl = torch.from_numpy(x[part.r0:part.r1].copy()).requires_grad_(True)
        y = b200dist.aggregate(part, xl, "wide")                 # two widths, one buffer key each
        z = b200dist.aggregate(part, (2 * y)[:, :5], "narrow")
        (z * torch.from_numpy(c[part.r0:part.r1])).sum().backward()
        assert {k[0] for k in part._full} == {"xwide", "dywide", "xnarrow", "dynarrow"}   # one buffer per (key, width, dtype)
        q.put((rank, z.detach().numpy(), xl.grad.numpy()))
    finally:
        dist.destroy_process_group()


def test_partition_autograd_two_widths_gloo():
    """dist.aggregate (what a row-partitioned GCN epoch calls twice per layer): forward values and
    the gradient through two chained aggregations of different widths, world_size 2."""
    world, n, e, feat = 2, 200, 3000, 8
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_autograd, args=(r, world, port, n, e, feat, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in range(world)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    src, dst = random_graph(n, e, 9, hubs=2, hub_deg=n // 3, zero_deg_frac=0.1)
    rng = np.random.RandomState(5)
    x = rng.randn(n, feat).astype(np.float32)
    c = rng.randn(n, 5).astype(np.float32)
    y = oracle.update_all(src, dst, n, "copy_src", "sum", x=x)
    z = oracle.update_all(src, dst, n, "copy_src", "sum", x=np.ascontiguousarray(2 * y[:, :5]))
    want_dx = np.zeros_like(x)
    want_dx[:, :5] = oracle.spmm_grad(src, dst, n, 2 * oracle.spmm_grad(src, dst, n, c))
    got_z = np.concatenate([r[1] for r in res])
    got_dx = np.concatenate([r[2] for r in res])
    np.testing.assert_allclose(got_z, z, rtol=1e-5, atol=1e-5 * np.abs(z).max())
    np.testing.assert_allclose(got_dx, want_dx, rtol=1e-5, atol=1e-5 * np.abs(want_dx).max())
