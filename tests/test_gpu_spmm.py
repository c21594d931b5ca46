"""-m gpu: kernels (1)/(2) SpMM vs the oracle through the C ABI.

Mirrors the reference's tests/compute/test_specialization.py (v2v / e2v / fallback
equivalence, 1-D / 2-D / 3-D features, scalar and (E,1) weights, zero-in-degree nodes,
fn.max) and tests/compute/test_function.py:24-49 (literal known answers).
Tolerance: fp32 rtol 1e-5 (north-star), bf16 rtol 1e-2; argmax ids bit-exact.
"""
import numpy as np
import pytest
import torch

from oracle import oracle
from tests.util import random_graph, hub_graph, HUB_EDGE_H

pytestmark = pytest.mark.gpu

RTOL_F32 = 1e-5
RTOL_BF16 = 1e-2


def _csr(src, dst, n, transpose=False):
    from dgl_b200 import ops
    s = torch.from_numpy(src).cuda()
    d = torch.from_numpy(dst).cuda()
    row, col = (s, d) if transpose else (d, s)
    return ops.csr_build(row, col, n, n, order=0)


def _close(got, want, rtol, scale=None):
    got = got.float().cpu().numpy()
    scale = np.abs(want).max() + 1e-30 if scale is None else scale
    np.testing.assert_allclose(got, want, rtol=rtol, atol=rtol * scale)


def test_known_answers_hub_graph():
    """tests/compute/test_function.py:24-49 of the reference."""
    from dgl_b200 import ops
    src, dst = hub_graph()
    csr = _csr(src, dst, 10)
    h = torch.arange(1, 11, dtype=torch.float32).cuda().reshape(10, 1)
    e = torch.tensor(HUB_EDGE_H, dtype=torch.float32).cuda().reshape(17, 1)
    y, _ = ops.spmm(csr, ops.MSG_COPY_SRC, ops.RED_SUM, x=h)
    assert y.flatten().tolist() == [10., 1., 1., 1., 1., 1., 1., 1., 1., 44.]
    y, _ = ops.spmm(csr, ops.MSG_COPY_EDGE, ops.RED_SUM, w=e)
    assert y.flatten().tolist() == [10., 1., 1., 1., 1., 1., 1., 1., 1., 44.]
    y, _ = ops.spmm(csr, ops.MSG_SRC_MUL_EDGE, ops.RED_SUM, x=h, w=e, w_width=1)
    assert y.flatten().tolist() == [100., 1., 1., 1., 1., 1., 1., 1., 1., 284.]


FEATS = [1, 3, 4, 5, 16, 41, 64, 128, 256, 260, 602]


@pytest.mark.parametrize("feat", FEATS)
@pytest.mark.parametrize("chunk", [0, 32, 512])
def test_copy_src_sum(feat, chunk):
    from dgl_b200 import ops
    n = 400
    src, dst = random_graph(n, 6000, feat, hubs=3, hub_deg=700, zero_deg_frac=0.2)
    csr = _csr(src, dst, n)
    x = np.random.RandomState(feat).randn(n, feat).astype(np.float32)
    want = oracle.update_all(src, dst, n, "copy_src", "sum", x=x)
    y, _ = ops.spmm(csr, ops.MSG_COPY_SRC, ops.RED_SUM, x=torch.from_numpy(x).cuda(), chunk=chunk)
    _close(y, want, RTOL_F32)
    if chunk == 0:
        # single-item rows add in edge-id order, the reference's COO order: bit-exact
        assert np.array_equal(y.cpu().numpy(), want)


@pytest.mark.parametrize("feat,wshape", [(1, ()), (8, ()), (8, (1,)), (64, ()), (64, (64,)),
                                         (64, (8,)), (60, (5,)), (7, (7,)), (256, ()),
                                         (24, (8,)), (256, (8,)), (6, (2,))])
@pytest.mark.parametrize("chunk", [32, 512])
def test_src_mul_edge_sum(feat, wshape, chunk):
    from dgl_b200 import ops
    n = 300
    src, dst = random_graph(n, 5000, 7 + feat, hubs=2, hub_deg=400)
    E = len(src)
    rng = np.random.RandomState(feat)
    x = rng.randn(n, feat).astype(np.float32)
    w = rng.rand(*((E,) + wshape)).astype(np.float32)
    csr = _csr(src, dst, n)
    w_width = int(np.prod(wshape)) if wshape else 1
    if w_width not in (1, feat):  # per-head layout (N,H,D) * (E,H,1)
        xs = x.reshape(n, w_width, feat // w_width)
        ws = w.reshape(E, w_width, 1)
    else:
        xs, ws = x, w
    want = oracle.update_all(src, dst, n, "src_mul_edge", "sum", x=xs, w=ws).reshape(n, feat)
    y, _ = ops.spmm(csr, ops.MSG_SRC_MUL_EDGE, ops.RED_SUM, x=torch.from_numpy(x).cuda(),
                    w=torch.from_numpy(w.reshape(E, -1)).cuda(), w_width=w_width, chunk=chunk)
    _close(y, want, RTOL_F32)


@pytest.mark.parametrize("feat", [1, 5, 8, 64, 200])
@pytest.mark.parametrize("reduce_op", ["sum", "max"])
def test_copy_edge(feat, reduce_op):
    from dgl_b200 import ops
    n = 200
    src, dst = random_graph(n, 3000, 11 + feat, hubs=2, hub_deg=300, zero_deg_frac=0.25)
    E = len(src)
    w = np.random.RandomState(feat).randn(E, feat).astype(np.float32)
    csr = _csr(src, dst, n)
    red = ops.RED_SUM if reduce_op == "sum" else ops.RED_MAX
    for chunk in (0, 64):
        y, arg = ops.spmm(csr, ops.MSG_COPY_EDGE, red, w=torch.from_numpy(w).cuda(),
                          want_arg=True, chunk=chunk)
        if reduce_op == "sum":
            _close(y, oracle.update_all(src, dst, n, "copy_edge", "sum", w=w), RTOL_F32)
        else:
            want, warg = oracle.reduce_max(dst, w, n)
            assert np.array_equal(y.cpu().numpy(), want)          # max is exact
            assert np.array_equal(arg.cpu().numpy().astype(np.int64), warg)


@pytest.mark.parametrize("feat", [1, 4, 6, 64, 256])
@pytest.mark.parametrize("msg", ["copy_src", "src_mul_edge"])
def test_max_with_argmax(feat, msg):
    from dgl_b200 import ops
    n = 150
    src, dst = random_graph(n, 2500, 3 + feat, hubs=2, hub_deg=500, zero_deg_frac=0.3)
    E = len(src)
    rng = np.random.RandomState(feat)
    # few distinct values => many ties: exercises the smallest-slot tie rule
    x = rng.randint(-3, 4, size=(n, feat)).astype(np.float32)
    w = rng.randint(1, 3, size=(E,)).astype(np.float32)
    csr = _csr(src, dst, n)
    m = oracle.message(msg, src, x, w if msg == "src_mul_edge" else None)
    want, warg = oracle.reduce_max(dst, m, n)
    for chunk in (0, 32, 512):
        y, arg = ops.spmm(csr, ops.MSG_COPY_SRC if msg == "copy_src" else ops.MSG_SRC_MUL_EDGE,
                          ops.RED_MAX, x=torch.from_numpy(x).cuda(),
                          w=torch.from_numpy(w).cuda() if msg == "src_mul_edge" else None,
                          w_width=1, want_arg=True, chunk=chunk)
        assert np.array_equal(y.cpu().numpy(), want)
        assert np.array_equal(arg.cpu().numpy().astype(np.int64), warg)


@pytest.mark.parametrize("feat", [8, 64, 256, 264])
def test_bf16(feat):
    from dgl_b200 import ops
    n = 300
    src, dst = random_graph(n, 4000, feat, hubs=2, hub_deg=600)
    x = torch.from_numpy(np.random.RandomState(1).randn(n, feat).astype(np.float32)).cuda().bfloat16()
    csr = _csr(src, dst, n)
    want = oracle.update_all(src, dst, n, "copy_src", "sum", x=x.float().cpu().numpy())
    y, _ = ops.spmm(csr, ops.MSG_COPY_SRC, ops.RED_SUM, x=x)
    assert y.dtype == torch.bfloat16
    _close(y, want, RTOL_BF16)


@pytest.mark.parametrize("feat,w_width", [(16, 0), (16, 1), (64, 8), (5, 5), (256, 0)])
def test_backward_x_sum(feat, w_width):
    """dx = A^T dy (torch's sparse.mm backward in the reference)."""
    from dgl_b200 import ops
    n = 250
    src, dst = random_graph(n, 4000, 5 + feat, hubs=2, hub_deg=400)
    E = len(src)
    rng = np.random.RandomState(feat)
    dy = rng.randn(n, feat).astype(np.float32)
    w = rng.rand(E, w_width).astype(np.float32) if w_width else None
    csr_t = _csr(src, dst, n, transpose=True)
    if w is not None and w_width not in (1, feat):
        want = oracle.spmm_grad(src, dst, n, dy.reshape(n, w_width, -1), w.reshape(E, w_width, 1)).reshape(n, feat)
    else:
        want = oracle.spmm_grad(src, dst, n, dy, w)
    dx = ops.spmm_backward_x(csr_t, torch.from_numpy(dy).cuda(),
                             w=torch.from_numpy(w).cuda() if w is not None else None,
                             w_width=max(w_width, 1), chunk=64)
    _close(dx, want, RTOL_F32)


def test_backward_x_max():
    from dgl_b200 import ops
    n, feat = 120, 12
    src, dst = random_graph(n, 1500, 42, hubs=1, hub_deg=300, zero_deg_frac=0.2)
    rng = np.random.RandomState(0)
    x = rng.randn(n, feat).astype(np.float32)
    dy = rng.randn(n, feat).astype(np.float32)
    csr = _csr(src, dst, n)
    csr_t = _csr(src, dst, n, transpose=True)
    _, arg = ops.spmm(csr, ops.MSG_COPY_SRC, ops.RED_MAX, x=torch.from_numpy(x).cuda(), want_arg=True)
    dx = ops.spmm_backward_x(csr_t, torch.from_numpy(dy).cuda(), arg=arg, chunk=32)
    warg = oracle.reduce_max(dst, x[src], n)[1]
    want = np.zeros((n, feat), dtype=np.float32)
    for v in range(n):
        for f in range(feat):
            if warg[v, f] >= 0:
                want[src[warg[v, f]], f] += dy[v, f]
    _close(dx, want, RTOL_F32)


def test_determinism_and_large():
    """Reddit-like degree skew at a size the oracle still handles; run-to-run bit-stable."""
    from dgl_b200 import ops
    n, feat = 20000, 128
    src, dst = random_graph(n, 300000, 9, hubs=8, hub_deg=5000)
    x = np.random.RandomState(3).randn(n, feat).astype(np.float32)
    csr = _csr(src, dst, n)
    xt = torch.from_numpy(x).cuda()
    y1, _ = ops.spmm(csr, ops.MSG_COPY_SRC, ops.RED_SUM, x=xt)
    y2, _ = ops.spmm(csr, ops.MSG_COPY_SRC, ops.RED_SUM, x=xt)
    assert torch.equal(y1, y2)
    import scipy.sparse as sp
    A = sp.csr_matrix((np.ones(len(src), np.float32), (dst, src)), shape=(n, n))
    want = (A.astype(np.float64) @ x.astype(np.float64)).astype(np.float32)
    _close(y1, want, RTOL_F32)


@pytest.mark.parametrize("feat,slab", [(256, 64), (128, 32), (96, 24), (12, 5)])
def test_strided_column_blocks(feat, slab):
    """dglb200_spmm_csr_strided: x and y are column blocks of wider row-major matrices (the
    feature-slab pipeline of dgl_b200/dist.py); results equal the dense call bit for bit."""
    from dgl_b200 import ops
    n = 500
    src, dst = random_graph(n, 9000, 21 + feat, hubs=2, hub_deg=900, zero_deg_frac=0.1)
    csr = _csr(src, dst, n)
    x = torch.from_numpy(np.random.RandomState(feat).randn(n, feat).astype(np.float32)).cuda()
    dense, _ = ops.spmm(csr, ops.MSG_COPY_SRC, ops.RED_SUM, x=x)
    y = torch.full((n, feat), float("nan"), device="cuda")
    for s0 in range(0, feat, slab):
        s1 = min(feat, s0 + slab)
        ops.spmm(csr, ops.MSG_COPY_SRC, ops.RED_SUM, x=x[:, s0:s1], out=y[:, s0:s1])
    assert torch.equal(y, dense)
    want = oracle.update_all(src, dst, n, "copy_src", "sum", x=x.cpu().numpy())
    _close(y, want, RTOL_F32)


def test_gather_rows():
    from dgl_b200 import ops
    rng = np.random.RandomState(0)
    for shape in ((1000, 128), (1000, 7), (300, 4, 6), (50,)):
        x = torch.from_numpy(rng.randn(*shape).astype(np.float32)).cuda()
        idx = torch.from_numpy(rng.randint(0, shape[0], size=777)).cuda()
        assert torch.equal(ops.gather_rows(x, idx), x[idx])
    xb = torch.randn(64, 40, device="cuda").bfloat16()
    idx = torch.randint(0, 64, (100,), device="cuda")
    assert torch.equal(ops.gather_rows(xb, idx), xb[idx])


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_slot_ordered_scalar_weights(dtype, monkeypatch):
    """dglb200_edge_data_to_slots + eids == NULL: the large-graph form of SPMV_WITH_DATA.
    The permutation is bit-exact; forward and dx equal the edge-id-indexed form bit for bit."""
    from dgl_b200 import ops
    n, feat = 500, 64
    src, dst = random_graph(n, 9000, 11, hubs=2, hub_deg=900, zero_deg_frac=0.1)
    csr, csr_t = _csr(src, dst, n), _csr(src, dst, n, transpose=True)
    rs = np.random.RandomState(3)
    x = torch.from_numpy(rs.randn(n, feat).astype(np.float32)).cuda().to(dtype)
    w = torch.from_numpy(rs.rand(len(src), 1).astype(np.float32)).cuda().to(dtype)
    ws = ops.edge_data_to_slots(csr, w)
    assert torch.equal(ws.reshape(-1), w.reshape(-1)[csr.eids.long()])
    y0, _ = ops.spmm(csr, ops.MSG_SRC_MUL_EDGE, ops.RED_SUM, x=x, w=w, w_width=1)
    dx0 = ops.spmm_backward_x(csr_t, x, w=w, w_width=1)
    monkeypatch.setattr(ops, "SLOT_ORDER_MIN_NNZ", 0)
    y1, _ = ops.spmm(csr, ops.MSG_SRC_MUL_EDGE, ops.RED_SUM, x=x, w=w, w_width=1)
    dx1 = ops.spmm_backward_x(csr_t, x, w=w, w_width=1)
    assert torch.equal(y0, y1) and torch.equal(dx0, dx1)
    if dtype == torch.float32:
        want = oracle.update_all(src, dst, n, "src_mul_edge", "sum", x=x.cpu().numpy(), w=w.cpu().numpy())
        _close(y1, want, RTOL_F32)
    # MAX with argmax keeps the edge-id form (ids are an output)
    y2, a2 = ops.spmm(csr, ops.MSG_SRC_MUL_EDGE, ops.RED_MAX, x=x, w=w, w_width=1, want_arg=True)
    monkeypatch.setattr(ops, "SLOT_ORDER_MIN_NNZ", 1 << 22)
    y3, a3 = ops.spmm(csr, ops.MSG_SRC_MUL_EDGE, ops.RED_MAX, x=x, w=w, w_width=1, want_arg=True)
    assert torch.equal(y2, y3) and torch.equal(a2, a3)


@pytest.mark.parametrize("feat", [4, 6, 64])
def test_max_with_minus_inf_messages(feat):
    """Rows whose messages are all -inf keep -inf with the first edge as argmax; zero-degree rows
    stay 0 / -1; mixed rows pick the first finite maximum (long-row slices included)."""
    from dgl_b200 import ops
    n = 120
    src, dst = random_graph(n, 1500, 77 + feat, hubs=2, hub_deg=300, zero_deg_frac=0.2)
    rng = np.random.RandomState(feat)
    x = rng.randint(-2, 3, size=(n, feat)).astype(np.float32)
    x[rng.rand(n) < 0.6] = -np.inf          # most sources send -inf on every feature
    x[rng.rand(n, feat) < 0.2] = -np.inf
    want, warg = oracle.reduce_max(dst, oracle.message("copy_src", src, x), n)
    csr = _csr(src, dst, n)
    for chunk in (0, 32):
        y, arg = ops.spmm(csr, ops.MSG_COPY_SRC, ops.RED_MAX, x=torch.from_numpy(x).cuda(),
                          want_arg=True, chunk=chunk)
        assert np.array_equal(y.cpu().numpy(), want)
        assert np.array_equal(arg.cpu().numpy().astype(np.int64), warg)


@pytest.mark.parametrize("feat", [4, 64])
def test_max_propagates_nan(feat):
    """The reference's fn.max is torch.max over the mailbox (function/reducer.py:45-46 ->
    backend/pytorch/tensor.py:95-97), which PROPAGATES NaN: a destination that receives a NaN
    message reads NaN on that feature, whatever else arrives (long-row slices included)."""
    from dgl_b200 import ops
    n = 100
    src, dst = random_graph(n, 1200, 5 + feat, hubs=2, hub_deg=200, zero_deg_frac=0.2)
    rng = np.random.RandomState(feat)
    x = rng.randn(n, feat).astype(np.float32)
    x[rng.rand(n, feat) < 0.02] = np.nan
    msgs = oracle.message("copy_src", src, x)
    want, _ = oracle.reduce_max(dst, msgs, n)
    # the reference value, stated with torch itself on a few rows
    for v in np.unique(dst)[:10]:
        ref = torch.max(torch.from_numpy(msgs[dst == v]), 0)[0].numpy()
        assert np.array_equal(ref, want[v], equal_nan=True)
    assert np.isnan(want).any()
    csr = _csr(src, dst, n)
    for chunk in (0, 32):
        y, arg = ops.spmm(csr, ops.MSG_COPY_SRC, ops.RED_MAX, x=torch.from_numpy(x).cuda(),
                          want_arg=True, chunk=chunk)
        assert np.array_equal(y.cpu().numpy(), want, equal_nan=True)


@pytest.mark.parametrize("feat", [41, 602, 90])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_ragged_width_takes_padded_vector_path(feat, dtype):
    """feat not a multiple of the 16-byte vector: ops.spmm pads the gathered rows (x_ld) and the
    kernel stores the ragged last chunk element-wise; a strided out block next to live data must
    not be clobbered."""
    from dgl_b200 import ops
    n = 300
    src, dst = random_graph(n, 5000, feat, hubs=2, hub_deg=600, zero_deg_frac=0.1)
    rs = np.random.RandomState(feat)
    x = rs.randn(n, feat).astype(np.float32)
    w = rs.rand(len(src), 1).astype(np.float32)
    csr = _csr(src, dst, n)
    xt, wt = torch.from_numpy(x).cuda().to(dtype), torch.from_numpy(w).cuda().to(dtype)
    rtol = RTOL_F32 if dtype == torch.float32 else RTOL_BF16
    xr, wr = xt.float().cpu().numpy(), wt.float().cpu().numpy()
    for msg, code, ww in (("copy_src", ops.MSG_COPY_SRC, None), ("src_mul_edge", ops.MSG_SRC_MUL_EDGE, wt)):
        want = oracle.update_all(src, dst, n, msg, "sum", x=xr, w=wr if ww is not None else None)
        for chunk in (0, 64):
            y, _ = ops.spmm(csr, code, ops.RED_SUM, x=xt, w=ww, w_width=1, chunk=chunk)
            _close(y, want, rtol)
        big = torch.full((n, feat + 7), 7.0, dtype=dtype, device="cuda")
        ops.spmm(csr, code, ops.RED_SUM, x=xt, w=ww, w_width=1, out=big[:, 3:3 + feat])
        _close(big[:, 3:3 + feat], want, rtol)
        assert bool((big[:, :3] == 7).all()) and bool((big[:, 3 + feat:] == 7).all())
    if dtype == torch.float32:
        m = oracle.message("copy_src", src, x)
        want, warg = oracle.reduce_max(dst, m, n)
        y, arg = ops.spmm(csr, ops.MSG_COPY_SRC, ops.RED_MAX, x=xt, want_arg=True, chunk=64)
        assert np.array_equal(y.cpu().numpy(), want)
        assert np.array_equal(arg.cpu().numpy().astype(np.int64), warg)
        dx = ops.spmm_backward_x(_csr(src, dst, n, transpose=True), xt, w=wt, w_width=1)
        _close(dx, oracle.spmm_grad(src, dst, n, x, w), rtol)


@pytest.mark.parametrize("heads,dim", [(8, 16), (4, 32), (8, 4), (2, 64)])
def test_per_head_weight_routes_agree_bitwise(heads, dim):
    """Per-head edge weights (E, H) as scalar-weight passes of one head (dglb200_set_tuning "spmm_head_pass" 1) and as
    chunk weights over the usual slabs (2): same products in the same order, same bits; both equal the oracle."""
    from dgl_b200 import _lib, ops
    n, feat = 240, heads * dim
    src, dst = random_graph(n, 3200, heads * 5 + dim, hubs=2, hub_deg=600, zero_deg_frac=0.2)
    rng = np.random.RandomState(dim)
    x = rng.randn(n, heads, dim).astype(np.float32)
    w = rng.rand(len(src), heads, 1).astype(np.float32)
    csr = _csr(src, dst, n)
    xt = torch.from_numpy(x.reshape(n, feat)).cuda()
    wt = torch.from_numpy(w.reshape(len(src), heads)).cuda()
    lib = _lib.get_lib()
    got = []
    try:
        for mode in (1, 2, 0):
            _lib.check_call(lib.dglb200_set_tuning(b"spmm_head_pass", mode))
            got.append(ops.spmm(csr, ops.MSG_SRC_MUL_EDGE, ops.RED_SUM, x=xt, w=wt, w_width=heads, chunk=32)[0])
    finally:
        _lib.check_call(lib.dglb200_set_tuning(b"spmm_head_pass", 0))
    assert torch.equal(got[0], got[1]) and torch.equal(got[0], got[2])
    want = oracle.update_all(src, dst, n, "src_mul_edge", "sum", x=x, w=w).reshape(n, feat)
    np.testing.assert_allclose(got[0].cpu().numpy(), want, rtol=1e-5, atol=1e-5 * np.abs(want).max())


@pytest.mark.parametrize("heads,dim", [(8, 8), (4, 32), (8, 3), (2, 4)])
def test_multi_head_weights_max_bf16_gradient(heads, dim):
    """(E, H, 1) weights against (N, H, D) features - the multi-head attention case
    (python/dgl/function/message.py:75-91 rank alignment): max + argmax, bf16 storage and the
    gradient wrt x, through the per-head chunk variant (D a multiple of the vector) and the
    generic one (D = 3)."""
    from dgl_b200 import ops
    n, feat = 260, heads * dim
    src, dst = random_graph(n, 4000, 13 + feat, hubs=2, hub_deg=600, zero_deg_frac=0.15)
    E = len(src)
    rng = np.random.RandomState(feat)
    x = rng.randint(-3, 4, size=(n, feat)).astype(np.float32)
    w = rng.randint(1, 4, size=(E, heads)).astype(np.float32)
    csr, csr_t = _csr(src, dst, n), _csr(src, dst, n, transpose=True)
    m = oracle.message("src_mul_edge", src, x.reshape(n, heads, dim), w.reshape(E, heads, 1)).reshape(E, feat)
    xt, wt = torch.from_numpy(x).cuda(), torch.from_numpy(w).cuda()
    want, warg = oracle.reduce_max(dst, m, n)
    for chunk in (0, 64):
        y, arg = ops.spmm(csr, ops.MSG_SRC_MUL_EDGE, ops.RED_MAX, x=xt, w=wt, w_width=heads,
                          want_arg=True, chunk=chunk)
        assert np.array_equal(y.cpu().numpy(), want)
        assert np.array_equal(arg.cpu().numpy().astype(np.int64), warg)
    # small integers are exact in bf16: the bf16 sum must equal the fp32 oracle after rounding
    ys, _ = ops.spmm(csr, ops.MSG_SRC_MUL_EDGE, ops.RED_SUM, x=xt.bfloat16(), w=wt.bfloat16(), w_width=heads)
    _close(ys, oracle.reduce_sum(dst, m, n), RTOL_BF16)
    dy = rng.randn(n, feat).astype(np.float32)
    want_dx = oracle.spmm_grad(src, dst, n, dy.reshape(n, heads, dim), w.reshape(E, heads, 1)).reshape(n, feat)
    dx = ops.spmm_backward_x(csr_t, torch.from_numpy(dy).cuda(), w=wt, w_width=heads)
    _close(dx, want_dx, RTOL_F32)
    # large-graph form: per-head weights permuted to slot order, eids == NULL; same bits
    y0, _ = ops.spmm(csr, ops.MSG_SRC_MUL_EDGE, ops.RED_SUM, x=xt, w=wt, w_width=heads)
    assert torch.equal(ops.edge_data_to_slots(csr, wt), wt[csr.eids.long()])
    assert torch.equal(ops.edge_data_to_slots(csr, wt.bfloat16()), wt.bfloat16()[csr.eids.long()])
    old = ops.SLOT_ORDER_MIN_NNZ
    ops.SLOT_ORDER_MIN_NNZ = 0
    try:
        y1, _ = ops.spmm(csr, ops.MSG_SRC_MUL_EDGE, ops.RED_SUM, x=xt, w=wt, w_width=heads)
        dx1 = ops.spmm_backward_x(csr_t, torch.from_numpy(dy).cuda(), w=wt, w_width=heads)
    finally:
        ops.SLOT_ORDER_MIN_NNZ = old
    assert torch.equal(y0, y1) and torch.equal(dx, dx1)


@pytest.mark.parametrize("feat", [8, 64, 128, 256])
@pytest.mark.parametrize("chunk", [0, 64])
def test_accumulate_and_bf16_halo_operand(feat, chunk):
    """dglb200_spmm_csr_ext, the two extensions of the multi-GPU row partition: (a) accumulate - the sum
    over two column blocks of the operand, second call adding into the first's output, is bit for bit
    y_a + y_b and within 1e-5 of the oracle over all edges; (b) column ids >= halo_first read from a
    bfloat16 second operand: identical bits to the fp32 kernel over the same rows widened beforehand."""
    from dgl_b200 import ops
    n, own = 400, 250
    src, dst = random_graph(n, 7000, 31 + feat, hubs=2, hub_deg=900, zero_deg_frac=0.2)
    rng = np.random.RandomState(feat)
    x = rng.randn(n, feat).astype(np.float32)
    xt = torch.from_numpy(x).cuda()
    # (a) split the edges by source: [0, own) and [own, n)
    local = src < own
    csr_a = _csr(src[local], dst[local], n)
    csr_b = _csr(src[~local], dst[~local], n)
    ya = ops.spmm(csr_a, ops.MSG_COPY_SRC, ops.RED_SUM, x=xt, chunk=chunk)[0]
    yb = ops.spmm(csr_b, ops.MSG_COPY_SRC, ops.RED_SUM, x=xt, chunk=chunk)[0]
    acc = ya.clone()
    ops.spmm(csr_b, ops.MSG_COPY_SRC, ops.RED_SUM, x=xt, chunk=chunk, out=acc, accumulate=True)
    assert torch.equal(acc, ya + yb)
    _close(acc, oracle.update_all(src, dst, n, "copy_src", "sum", x=x), RTOL_F32)
    # a strided column block of a wider output, accumulated in place
    wide = torch.zeros(n, feat + 8, device="cuda")
    wide[:, 4:4 + feat] = ya
    if feat % 4 == 0:
        ops.spmm(csr_b, ops.MSG_COPY_SRC, ops.RED_SUM, x=xt, chunk=chunk, out=wide[:, 4:4 + feat], accumulate=True)
        assert torch.equal(wide[:, 4:4 + feat], ya + yb) and float(wide[:, :4].abs().max()) == 0.0
    with pytest.raises(Exception):
        ops.spmm(csr_b, ops.MSG_COPY_SRC, ops.RED_MAX, x=xt, out=acc, accumulate=True)
    # (b) rows [own, n) of the operand live in bf16
    csr = _csr(src, dst, n)
    halo = xt[own:].bfloat16().contiguous()
    widened = torch.cat([xt[:own], halo.float()])
    want = ops.spmm(csr, ops.MSG_COPY_SRC, ops.RED_SUM, x=widened, chunk=chunk)[0]
    got = ops.spmm(csr, ops.MSG_COPY_SRC, ops.RED_SUM, x=xt[:own].contiguous(), chunk=chunk, x_halo=halo,
                   halo_first=own)[0]
    assert torch.equal(got, want)
    _close(got, oracle.update_all(src, dst, n, "copy_src", "sum", x=x), RTOL_BF16)
