"""-m gpu: kernels (3) SDDMM / edge softmax / fused GAT and (4b) degree bucketing / unique vs
the oracle.  GAT mirrors examples/pytorch/gat/train.py:61-97; bucketing the canonical form of
src/scheduler/scheduler.cc:13-93 (shape sets of tests/compute/test_basics.py:324-361)."""
import numpy as np
import pytest
import torch

from oracle import oracle
from tests.util import random_graph, hub_graph

pytestmark = pytest.mark.gpu


def _csr(src, dst, n):
    from dgl_b200 import ops
    return ops.csr_build(torch.from_numpy(dst).cuda(), torch.from_numpy(src).cuda(), n, n)


@pytest.mark.parametrize("heads,dim", [(1, 1), (8, 1), (8, 8), (1, 256), (2, 100), (3, 5), (4, 32),
                                       # shapes of the warp-per-slot-run DOT kernel: a head = 2..32 chunks, idle lanes,
                                       # whole-warp heads (dim 256 / 512), 8 chunk registers per lane, and its fallbacks
                                       (2, 128), (3, 8), (1, 512), (8, 128), (16, 64), (5, 24), (2, 1024)])
@pytest.mark.parametrize("op", ["dot", "add", "mul"])
def test_sddmm(heads, dim, op):
    from dgl_b200 import ops
    n = 300
    src, dst = random_graph(n, 4000, heads * 31 + dim, hubs=2, hub_deg=500, zero_deg_frac=0.2)
    rng = np.random.RandomState(dim)
    u = rng.randn(n, heads, dim).astype(np.float32)
    v = rng.randn(n, heads, dim).astype(np.float32)
    want = oracle.sddmm(src, dst, u, v, op).reshape(len(src), -1)
    code = dict(dot=ops.SDDMM_DOT, add=ops.SDDMM_ADD, mul=ops.SDDMM_MUL)[op]
    got = ops.sddmm(_csr(src, dst, n), code, torch.from_numpy(u).cuda().reshape(n, -1),
                    torch.from_numpy(v).cuda().reshape(n, -1), heads, dim)
    np.testing.assert_allclose(got.cpu().numpy(), want, rtol=1e-5, atol=1e-5 * np.abs(want).max())


@pytest.mark.parametrize("heads,dim", [(8, 32), (2, 8), (1, 256), (3, 8), (8, 128), (1, 512)])
@pytest.mark.parametrize("dtype", ["f32", "bf16"])
def test_sddmm_dot_warp_kernel_equals_generic_bitwise(heads, dim, dtype):
    """The warp-per-slot-run DOT keeps the generic kernel's lane-local sums and butterfly order: same bits."""
    from dgl_b200 import _lib, ops
    n = 220
    src, dst = random_graph(n, 3000, heads * 7 + dim, hubs=2, hub_deg=400, zero_deg_frac=0.2)
    rng = np.random.RandomState(dim)
    u = torch.from_numpy(rng.randn(n, heads * dim).astype(np.float32)).cuda()
    v = torch.from_numpy(rng.randn(n, heads * dim).astype(np.float32)).cuda()
    if dtype == "bf16":
        u, v = u.bfloat16(), v.bfloat16()
    csr = _csr(src, dst, n)
    lib = _lib.get_lib()
    fast = ops.sddmm(csr, ops.SDDMM_DOT, u, v, heads, dim)
    try:
        _lib.check_call(lib.dglb200_set_tuning(b"sddmm_dot_generic", 1))
        slow = ops.sddmm(csr, ops.SDDMM_DOT, u, v, heads, dim)
    finally:
        _lib.check_call(lib.dglb200_set_tuning(b"sddmm_dot_generic", 0))
    assert torch.equal(fast, slow)


@pytest.mark.parametrize("heads,dim", [(8, 32), (2, 8), (1, 256), (3, 40)])
def test_sddmm_dot_bf16(heads, dim):
    """bf16 storage, fp32 accumulation (8-element chunks in the warp kernel; (3, 40) takes the generic one): against
    the oracle on the bf16-rounded operands, 1e-2 bar of the north-star for the stored result."""
    from dgl_b200 import ops
    n = 200
    src, dst = random_graph(n, 2500, heads + dim, hubs=1, hub_deg=300, zero_deg_frac=0.1)
    rng = np.random.RandomState(heads)
    u = torch.from_numpy(rng.randn(n, heads * dim).astype(np.float32)).cuda().bfloat16()
    v = torch.from_numpy(rng.randn(n, heads * dim).astype(np.float32)).cuda().bfloat16()
    got = ops.sddmm(_csr(src, dst, n), ops.SDDMM_DOT, u, v, heads, dim)
    assert got.dtype == torch.bfloat16
    want = oracle.sddmm(src, dst, u.float().cpu().numpy().reshape(n, heads, dim),
                        v.float().cpu().numpy().reshape(n, heads, dim), "dot").reshape(len(src), -1)
    np.testing.assert_allclose(got.float().cpu().numpy(), want, rtol=1e-2, atol=1e-2 * np.abs(want).max())


@pytest.mark.parametrize("heads", [1, 3, 8])
def test_edge_softmax_fwd_bwd(heads):
    from dgl_b200 import ops
    n = 200
    src, dst = random_graph(n, 3000, heads, hubs=2, hub_deg=400, zero_deg_frac=0.2)
    E = len(src)
    rng = np.random.RandomState(heads)
    s = (rng.randn(E, heads) * 3).astype(np.float32)
    csr = _csr(src, dst, n)
    alpha = ops.edge_softmax(csr, torch.from_numpy(s).cuda())
    want = oracle.edge_softmax(dst, s, n)
    np.testing.assert_allclose(alpha.cpu().numpy(), want, rtol=1e-5, atol=1e-6)
    # backward against torch autograd of the same formula (fp32 reference of the op)
    st = torch.from_numpy(s).cuda().requires_grad_()
    d = torch.from_numpy(dst).cuda()
    mx = torch.full((n, heads), -float("inf"), device="cuda").scatter_reduce(0, d[:, None].expand(-1, heads), st, "amax")
    ex = torch.exp(st - mx[d])
    den = torch.zeros(n, heads, device="cuda").index_add(0, d, ex)
    a_ref = ex / den[d]
    g = torch.from_numpy(rng.randn(E, heads).astype(np.float32)).cuda()
    a_ref.backward(g)
    ds = ops.edge_softmax_backward(csr, alpha, g)
    np.testing.assert_allclose(ds.cpu().numpy(), st.grad.cpu().numpy(), rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("heads,dim", [(8, 8), (1, 16), (2, 5), (8, 32), (4, 100)])
@pytest.mark.parametrize("flavour", ["example", "stable"])
def test_gat_aggregate(heads, dim, flavour):
    from dgl_b200 import ops
    n = 250
    src, dst = random_graph(n, 3000, heads + dim, hubs=2, hub_deg=300)
    loops = np.arange(n, dtype=np.int64)     # the example adds self loops (train.py:186-189)
    src, dst = np.concatenate([src, loops]), np.concatenate([dst, loops])
    rng = np.random.RandomState(dim)
    a1 = rng.randn(n, heads, 1).astype(np.float32)
    a2 = rng.randn(n, heads, 1).astype(np.float32)
    ft = rng.randn(n, heads, dim).astype(np.float32)
    csr = _csr(src, dst, n)
    fl = ops.SOFTMAX_GAT_EXAMPLE if flavour == "example" else ops.SOFTMAX_STABLE
    out, a, z = ops.gat_aggregate(csr, torch.from_numpy(a1).cuda(), torch.from_numpy(a2).cuda(),
                                  torch.from_numpy(ft).cuda(), 0.2, fl)
    if flavour == "example":
        w_out, w_a, w_z = oracle.gat_example(src, dst, n, a1, a2, ft, 0.2)
        np.testing.assert_allclose(a.cpu().numpy(), w_a.reshape(len(src), heads), rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(z.cpu().numpy(), w_z.reshape(n, heads), rtol=1e-5, atol=1e-5)
    else:
        s = oracle.leaky_relu(a1[src] + a2[dst], 0.2).reshape(len(src), heads)
        alpha = oracle.edge_softmax(dst, s, n)
        w_out = oracle.reduce_sum(dst, ft[src] * alpha[:, :, None], n)
        np.testing.assert_allclose(a.cpu().numpy(), alpha, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(out.cpu().numpy(), w_out, rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("heads,dim,slab", [(8, 32, 2), (8, 8, 3), (5, 4, 1), (4, 100, 2), (3, 5, 2), (8, 32, 8)])
@pytest.mark.parametrize("flavour", ["example", "stable"])
def test_gat_aggregate_head_slabs_bitwise(heads, dim, slab, flavour):
    """The fused kernel in passes of `slab` heads (what it does by itself once the gathered operand outgrows the L2;
    forced here through dglb200_set_tuning): output, per-edge attention and z bit-identical to the single pass -
    heads are independent and a row's accumulation order does not change.  (3, 5, 2) / (4, 100, 2): the scalar and the
    vector path with a ragged last slab."""
    from dgl_b200 import _lib, ops
    n = 240
    src, dst = random_graph(n, 2800, heads * 3 + dim, hubs=2, hub_deg=300)
    loops = np.arange(n, dtype=np.int64)
    src, dst = np.concatenate([src, loops]), np.concatenate([dst, loops])
    rng = np.random.RandomState(slab)
    a1 = torch.from_numpy(rng.randn(n, heads).astype(np.float32)).cuda()
    a2 = torch.from_numpy(rng.randn(n, heads).astype(np.float32)).cuda()
    ft = torch.from_numpy(rng.randn(n, heads, dim).astype(np.float32)).cuda()
    csr = _csr(src, dst, n)
    fl = ops.SOFTMAX_GAT_EXAMPLE if flavour == "example" else ops.SOFTMAX_STABLE
    lib = _lib.get_lib()
    one = ops.gat_aggregate(csr, a1, a2, ft, 0.2, fl)
    l0 = _lib.launch_count()
    try:
        _lib.check_call(lib.dglb200_set_tuning(b"gat_head_slab", slab))
        got = ops.gat_aggregate(csr, a1, a2, ft, 0.2, fl)
        launches = _lib.launch_count() - l0
    finally:
        _lib.check_call(lib.dglb200_set_tuning(b"gat_head_slab", 0))
    assert launches == -(-heads // slab)                       # one launch per head slab
    for a, b in zip(got, one):
        assert torch.equal(a, b)


BUCKET_CASES = [(10, 17, 0, 0), (50, 400, 1, 10), (1000, 20000, 2, 300), (5, 0, 3, 5),
                (30000, 200000, 4, 1000), (7, 3000, 5, 0)]


@pytest.mark.parametrize("n,m,seed,extra_recv", BUCKET_CASES)
def test_degree_bucketing(n, m, seed, extra_recv):
    from dgl_b200 import ops
    rng = np.random.RandomState(seed)
    vids = rng.randint(0, max(1, n // 2), size=m).astype(np.int64)
    msg_ids = rng.permutation(m).astype(np.int64)
    recv = np.unique(np.concatenate([vids, rng.randint(0, n, size=extra_recv)])).astype(np.int64)
    rng.shuffle(recv)
    want = oracle.degree_bucketing(msg_ids, vids, recv)
    got = ops.degree_bucketing(torch.from_numpy(msg_ids).cuda(), torch.from_numpy(vids).cuda(),
                               torch.from_numpy(recv).cuda(), n)
    for g, w, name in zip(got, want, ["degs", "nids", "nid_section", "mids", "mid_section"]):
        assert np.array_equal(g.cpu().numpy(), w), name


def test_degree_bucketing_hub_graph_shapes():
    """update_all on the hub graph buckets as {deg 1: 9 nodes, deg 8: 1 node}
    (tests/compute/test_basics.py:324-361 asserts the shape set {(1,8,D),(9,1,D)})."""
    from dgl_b200 import ops
    src, dst = hub_graph()
    mids = torch.arange(17, device="cuda")
    degs, nids, nsec, m, msec = ops.degree_bucketing(mids, torch.from_numpy(dst).cuda(),
                                                     torch.arange(10, device="cuda"), 10)
    assert degs.tolist() == [1, 8] and nsec.tolist() == [9, 1] and msec.tolist() == [9, 8]
    assert nids.tolist() == [0, 1, 2, 3, 4, 5, 6, 7, 8, 9]


@pytest.mark.parametrize("n,k,seed", [(10, 30, 0), (100000, 500000, 1), (5, 1, 2), (70000, 100, 3)])
def test_unique_sorted(n, k, seed):
    from dgl_b200 import ops
    ids = np.random.RandomState(seed).randint(0, n, size=k).astype(np.int64)
    uniq, inv = ops.unique_sorted(torch.from_numpy(ids).cuda(), n)
    w_u, w_i = oracle.unique_sorted(ids)
    assert np.array_equal(uniq.cpu().numpy(), w_u)
    assert np.array_equal(inv.cpu().numpy(), w_i)
