"""-m "not gpu": pins the oracle (oracle/oracle.py numpy restatement, oracle/oracle.c plain-C
restatement) against
  (a) the literal known answers of the reference's tests/compute/test_function.py:24-49,
  (b) golden vectors produced by the UNMODIFIED reference (oracle/gen_golden.py -> tests/golden/),
  (c) where oracle/_ref/libdgl.so exists, the reference's own C++ GetAdj / DegreeBucketing called
      live through its C ABI (oracle/ref_ffi.py).
Tolerance for floating point: rtol = atol-scale 1e-5 (north-star); integer outputs bit-exact.
The reference's own tests use rtol=atol=1e-4 (tests/backend/pytorch/__init__.py:11-12).
"""
import glob
import os

import numpy as np
import pytest

from oracle import c_oracle, oracle, ref_ffi
from tests.util import hub_graph, HUB_EDGE_H, random_graph

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = sorted(glob.glob(os.path.join(GOLDEN, "case_*.npz")))
RTOL = 1e-5


def close(got, want, rtol=RTOL):
    want = np.asarray(want)
    got = np.asarray(got).reshape(want.shape)
    np.testing.assert_allclose(got, want, rtol=rtol, atol=rtol * (np.abs(want).max() + 1e-30))


@pytest.fixture(scope="module", params=CASES, ids=[os.path.basename(c) for c in CASES])
def case(request):
    return dict(np.load(request.param))


def test_golden_files_present():
    assert len(CASES) >= 3
    assert os.path.exists(os.path.join(GOLDEN, "hub.npz"))
    assert os.path.exists(os.path.join(GOLDEN, "plans.json"))


def test_known_answers_reference_test_function():
    """tests/compute/test_function.py:24-49 literals, via numpy and C oracles and the golden file."""
    src, dst = hub_graph()
    hub = np.load(os.path.join(GOLDEN, "hub.npz"))
    assert np.array_equal(hub["src"], src) and np.array_equal(hub["dst"], dst)
    h = np.arange(1, 11, dtype=np.float32)
    e = np.asarray(HUB_EDGE_H, dtype=np.float32)
    want_copy = [10., 1., 1., 1., 1., 1., 1., 1., 1., 44.]
    want_mul = [100., 1., 1., 1., 1., 1., 1., 1., 1., 284.]
    assert oracle.update_all(src, dst, 10, "copy_src", "sum", x=h).tolist() == want_copy
    assert oracle.update_all(src, dst, 10, "copy_edge", "sum", w=e).tolist() == want_copy
    assert oracle.update_all(src, dst, 10, "src_mul_edge", "sum", x=h, w=e).tolist() == want_mul
    assert hub["copy_src"].tolist() == want_copy and hub["src_mul_edge"].tolist() == want_mul
    if c_oracle.available():
        assert c_oracle.spmm_sum(src, dst, 10, "copy_src", x=h).ravel().tolist() == want_copy
        assert c_oracle.spmm_sum(src, dst, 10, "copy_edge", w=e).ravel().tolist() == want_copy
        assert c_oracle.spmm_sum(src, dst, 10, "src_mul_edge", x=h, w=e).ravel().tolist() == want_mul


def test_structure_bit_exact(case):
    n, src, dst = int(case["n"]), case["src"], case["dst"]
    for tr, tag in ((False, "adj_csr"), (True, "adj_csr_t")):
        got = oracle.get_adj_csr(src, dst, n, transpose=tr)
        for g, k in zip(got, ("_indptr", "_indices", "_eid")):
            assert np.array_equal(g, case[tag + k]), tag + k
        if c_oracle.available():
            for g, k in zip(c_oracle.get_adj_csr(src, dst, n, transpose=tr), ("_indptr", "_indices", "_eid")):
                assert np.array_equal(g, case[tag + k]), "C " + tag + k
    assert np.array_equal(oracle.in_degrees(dst, n), case["in_degrees"])
    assert np.array_equal(oracle.in_degrees(src, n), case["out_degrees"])
    # ImmutableGraph::CSR::FromEdges: std::sort leaves duplicate (src,dst) pairs' eid order
    # unspecified -> compare indptr / indices exactly and eids as per-(row, col) multisets
    ip, ix, ei = oracle.immutable_csr(src, dst, n, sort_on=1)
    assert np.array_equal(ip, case["imm_csr_indptr"])
    assert np.array_equal(ix, case["imm_csr_indices"])
    rows = np.repeat(np.arange(n), np.diff(ip))
    k1 = np.lexsort((ei, ix, rows))
    k2 = np.lexsort((case["imm_csr_eid"], case["imm_csr_indices"], rows))
    assert np.array_equal(ei[k1], case["imm_csr_eid"][k2])


PAIRS_MUL = [("f1", "ws"), ("f3", "ws"), ("f3", "w1"), ("f64", "ws"), ("f64", "w1"), ("f55", "ws"),
             ("f55", "w55"), ("f46", "w41"), ("f64", "w64")]


def _w_for_c(x, w):
    """Flatten a broadcastable weight to the (E, w_width) trailing-group form the kernels use."""
    return w.reshape(len(w), -1)


def test_update_all_all_builtin_pairs(case):
    n, src, dst = int(case["n"]), case["src"], case["dst"]
    for fk in ("f1", "f3", "f55", "f64", "f46"):
        x = case["in_" + fk]
        for red in ("sum", "max"):
            want = case["copy_src_%s_%s" % (red, fk)]
            close(oracle.update_all(src, dst, n, "copy_src", red, x=x), want)
            if c_oracle.available():
                got = (c_oracle.spmm_sum if red == "sum" else lambda *a, **k: c_oracle.spmm_max(*a, **k)[0])(
                    src, dst, n, "copy_src", x=x)
                close(got, want)
    for fk, wk in PAIRS_MUL:
        x, w = case["in_" + fk], case["in_" + wk]
        for red in ("sum", "max"):
            want = case["src_mul_edge_%s_%s_%s" % (red, fk, wk)]
            close(oracle.update_all(src, dst, n, "src_mul_edge", red, x=x, w=w), want)
            if c_oracle.available():
                got = (c_oracle.spmm_sum if red == "sum" else lambda *a, **k: c_oracle.spmm_max(*a, **k)[0])(
                    src, dst, n, "src_mul_edge", x=x, w=_w_for_c(x, w))
                close(got, want)
    for wk in ("ws", "w1", "w55", "w64"):
        w = case["in_" + wk]
        for red in ("sum", "max"):
            want = case["copy_edge_%s_%s" % (red, wk)]
            close(oracle.update_all(src, dst, n, "copy_edge", red, w=w), want)
            if c_oracle.available():
                got = (c_oracle.spmm_sum if red == "sum" else lambda *a, **k: c_oracle.spmm_max(*a, **k)[0])(
                    src, dst, n, "copy_edge", w=w)
                close(got, want)


def test_max_is_exact_and_zero_degree_rows_are_zero(case):
    n, src, dst = int(case["n"]), case["src"], case["dst"]
    x = case["in_f64"]
    got, arg = oracle.reduce_max(dst, x[src], n)
    assert np.array_equal(got, case["copy_src_max_f64"])
    assert (got[n - 8:] == 0).all() and (arg[n - 8:] == -1).all()
    if c_oracle.available():
        cg, ca = c_oracle.spmm_max(src, dst, n, "copy_src", x=x)
        assert np.array_equal(cg, got) and np.array_equal(ca, arg)


def test_gradients(case):
    n, src, dst = int(case["n"]), case["src"], case["dst"]
    dy, x = case["in_dy64"], case["in_f64"]
    close(oracle.spmm_grad(src, dst, n, dy), case["grad_copy_src_sum_f64_dx"])
    close(oracle.spmm_grad(src, dst, n, dy, case["in_ws"]), case["grad_src_mul_edge_sum_f64_ws_dx"])
    close(oracle.spmm_grad(src, dst, n, dy, case["in_w64"]), case["grad_src_mul_edge_sum_f64_w64_dx"])
    close(oracle.sddmm(src, dst, x, dy, "dot"), case["grad_src_mul_edge_sum_f64_ws_dw"])
    close(oracle.sddmm(src, dst, x, dy, "mul"), case["grad_src_mul_edge_sum_f64_w64_dw"])
    if c_oracle.available():
        close(c_oracle.spmm_grad_x(src, dst, n, dy), case["grad_copy_src_sum_f64_dx"])
        close(c_oracle.spmm_grad_x(src, dst, n, dy, case["in_ws"]), case["grad_src_mul_edge_sum_f64_ws_dx"])
    # max: gradient flows to the argmax source only
    _, arg = oracle.reduce_max(dst, x[src], n)
    want = np.zeros_like(x)
    rows, cols = np.nonzero(arg >= 0)
    np.add.at(want, (src[arg[rows, cols]], cols), dy[rows, cols])
    close(want, case["grad_copy_src_max_f64_dx"])


def test_subsets_send_and_recv_pull(case):
    """Unique-ascending output order + untouched rows keep their old values
    (python/dgl/runtime/scheduler.py:111-156, 309-361)."""
    n, src, dst = int(case["n"]), case["src"], case["dst"]
    x, w = case["in_f64"], case["in_ws"]
    eids, pull_nodes = case["in_snr_eids"], case["in_pull_nodes"]
    for name in ("copy_src", "src_mul_edge"):
        for red in ("sum", "max"):
            recv, inv = oracle.unique_sorted(dst[eids])
            y = oracle.update_all(src[eids], inv, len(recv), name, red, x=x, w=w[eids])
            want = x.copy()
            want[recv] = y
            close(want, case["snr_%s_%s_h" % (name, red)])
            nodes = np.unique(pull_nodes)
            sel = np.nonzero(np.isin(dst, nodes))[0]
            order = np.argsort(np.searchsorted(nodes, dst[sel]), kind="stable")
            sel = sel[order]
            y = oracle.update_all(src[sel], np.searchsorted(nodes, dst[sel]), len(nodes), name, red,
                                  x=x, w=w[sel])
            want = x.copy()
            want[nodes] = y
            close(want, case["pull_%s_%s_h" % (name, red)])


def test_gat_example_edge_path(case):
    n, src, dst = int(case["n"]), case["src"], case["dst"]
    s2 = np.concatenate([src, np.arange(n)])
    d2 = np.concatenate([dst, np.arange(n)])
    out, a, z = oracle.gat_example(s2, d2, n, case["in_gat_a1"], case["in_gat_a2"], case["in_gat_ft"], 0.2)
    close(a, case["gat_a"])
    close(z, case["gat_z"])
    close(out, case["gat_out"])
    # the stable softmax gives the same normalised attention whenever exp does not clamp
    s = oracle.leaky_relu(case["in_gat_a1"][s2] + case["in_gat_a2"][d2], 0.2)
    if np.exp(s).max() < 10:
        alpha = oracle.edge_softmax(d2, s.reshape(len(s2), -1), n)
        close(alpha, (a / z[d2]).reshape(len(s2), -1), rtol=1e-4)
        if c_oracle.available():
            close(c_oracle.edge_softmax(d2, s.reshape(len(s2), -1), n), alpha)


def test_degree_bucketing_canonical_form(case):
    n, dst = int(case["n"]), case["dst"]
    e = len(dst)
    ref = oracle.canonicalize_bucketing(*[case["bucket_full_%d" % i] for i in range(5)])
    got = oracle.degree_bucketing(np.arange(e), dst, case["in_bucket_full_recv"])
    for g, r in zip(got, ref):
        assert np.array_equal(g, r)
    sub, recv = case["in_bucket_sub_mids"], case["in_bucket_sub_recv"]
    ref = oracle.canonicalize_bucketing(*[case["bucket_sub_%d" % i] for i in range(5)])
    got = oracle.degree_bucketing(sub, dst[sub], recv)
    for g, r in zip(got, ref):
        assert np.array_equal(g, r)
    if c_oracle.available():
        for g, r in zip(c_oracle.degree_bucketing(sub, dst[sub], recv, n), ref):
            assert np.array_equal(g, r)


@pytest.mark.skipif(not ref_ffi.available(), reason="oracle/_ref/libdgl.so not built")
@pytest.mark.parametrize("seed", [0, 1, 2, 3])
def test_live_reference_cpp(seed):
    """The compiled reference C++ itself, through its C ABI, on fresh seeded inputs."""
    n = 300
    src, dst = random_graph(n, 4000, seed, hubs=2, hub_deg=300, zero_deg_frac=0.2)
    g = ref_ffi.RefGraph(src, dst, n)
    for tr in (False, True):
        for a, b in zip(g.get_adj_csr(tr), oracle.get_adj_csr(src, dst, n, tr)):
            assert np.array_equal(a, b)
    idx, eid = g.get_adj_coo(False)
    want_idx, want_eid = oracle.get_adj_coo(src, dst)
    assert np.array_equal(idx, want_idx) and np.array_equal(eid, want_eid)
    assert np.array_equal(g.in_degrees(np.arange(n)), oracle.in_degrees(dst, n))
    gi = ref_ffi.RefGraph(src, dst, n, readonly=True)
    ip, ix, _ = gi.get_adj_csr(False)
    wip, wix, _ = oracle.immutable_csr(src, dst, n, sort_on=1)
    assert np.array_equal(ip, wip) and np.array_equal(ix, wix)
    recv = np.unique(np.concatenate([dst, np.arange(n - 5, n)]))
    raw = ref_ffi.degree_bucketing(np.arange(len(dst)), dst, recv)
    for a, b in zip(oracle.canonicalize_bucketing(*raw), oracle.degree_bucketing(np.arange(len(dst)), dst, recv)):
        assert np.array_equal(a, b)
