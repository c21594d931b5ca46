"""-m gpu, collected LAST on purpose: dglb200_spmm_csr_static_scratch, the
allocation-free SpMM entry point added (no counterpart in the
reference).  Its long-row scratch is a file-scope `static __device__` array of DGLB200_STATIC_SCRATCH_BYTES
inside the library instead of a caller-provided buffer.

Called straight through the C ABI (ctypes) and compared with the existing entry point dglb200_spmm_csr on
the same inputs, two calls in a row (the scratch is shared and keeps the previous call's contents), with
hub rows so that the long-row slices - the only users of the scratch - really run.  Same kernels, same
slice order: the results are expected bit-identical; the assertion uses the suite's fp32 bar (rtol 1e-5).
Self-contained: passes alone in a fresh process and after every other test.
"""
import ctypes

import numpy as np
import pytest
import torch

from tests.util import random_graph

pytestmark = pytest.mark.gpu
RTOL = 1e-5
SCRATCH_BYTES = 4 << 20          # DGLB200_STATIC_SCRATCH_BYTES (include/dgl_b200.h)


def device():
    return torch.device("cuda")


def _vp(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def _stream():
    from dgl_b200 import ops
    return ops._stream()


def _plan_struct(csr, chunk, feat, with_partial):
    """dglb200_spmm_plan_t over the caller's tables; partial only for the existing entry point."""
    from dgl_b200 import _lib
    plan = csr.plan(chunk)
    assert plan.n_long_rows > 0, "the test graph must have long rows, or no scratch is exercised"
    st = _lib.SpmmPlanStruct()
    st.chunk, st.n_long_rows, st.n_chunks = plan.chunk, plan.n_long_rows, plan.n_chunks
    st.long_rows, st.chunk_ptr = plan.long_rows.data_ptr(), plan.chunk_ptr.data_ptr()
    keep = None
    if with_partial:
        keep = torch.empty(plan.n_chunks * feat, dtype=torch.float32, device=csr.device)
        st.partial, st.partial_bytes = keep.data_ptr(), keep.numel() * 4
    else:
        st.partial, st.partial_bytes = None, 0          # ignored by the static-scratch entry point
    st.partial_arg = None
    return st, plan, keep


def _call(fn_name, csr, msg_op, red_op, x, w, w_width, y, st):
    from dgl_b200 import _lib
    lib = _lib.get_lib()
    feat = int(y.shape[1])
    args = [csr.n_rows, csr.n_cols, csr.nnz, _vp(csr.indptr), _vp(csr.indices), _vp(csr.eids), msg_op, red_op, 0,
            _vp(x), feat, _vp(w), w_width, _vp(y)]
    if fn_name == "dglb200_spmm_csr":
        args.append(ctypes.c_void_p(0))                   # arg (argmax ids): none
    args += [ctypes.byref(st), _stream()]
    return getattr(lib, fn_name)(*args)


def test_static_scratch_spmm_equals_existing_entry_point_twice_in_a_row():
    from dgl_b200 import _lib, ops
    dev = device()
    n, feat, chunk = 600, 64, 32
    src, dst = random_graph(n, 9000, 123, hubs=4, hub_deg=1500, zero_deg_frac=0.2)
    csr = ops.csr_build(torch.from_numpy(dst).to(dev), torch.from_numpy(src).to(dev), n, n)
    rng = np.random.RandomState(5)
    w = torch.from_numpy(rng.rand(len(src), 1).astype(np.float32)).to(dev)
    st_ref, plan, keep = _plan_struct(csr, chunk, feat, with_partial=True)
    st_new, _, _ = _plan_struct(csr, chunk, feat, with_partial=False)
    assert plan.n_chunks * feat * 4 <= SCRATCH_BYTES
    launches0 = _lib.launch_count()
    # two calls in a row on the shared scratch, different operands and ops: the second must not see the first
    for call, (msg_op, red_op, scale) in enumerate([(ops.MSG_COPY_SRC, ops.RED_SUM, 1.0),
                                                    (ops.MSG_SRC_MUL_EDGE, ops.RED_SUM, -3.0),
                                                    (ops.MSG_COPY_SRC, ops.RED_MAX, 0.5)]):
        x = torch.from_numpy((rng.randn(n, feat) * scale).astype(np.float32)).to(dev)
        ww = w if msg_op == ops.MSG_SRC_MUL_EDGE else None
        want = torch.full((n, feat), float("nan"), device=dev)
        got = torch.full((n, feat), float("nan"), device=dev)
        assert _call("dglb200_spmm_csr", csr, msg_op, red_op, x, ww, 1, want, st_ref) == 0, _lib.last_error()
        assert _call("dglb200_spmm_csr_static_scratch", csr, msg_op, red_op, x, ww, 1, got, st_new) == 0, _lib.last_error()
        torch.cuda.synchronize() if dev.type == "cuda" else None
        a, b = got.cpu().numpy(), want.cpu().numpy()
        assert not np.isnan(a).any()
        np.testing.assert_allclose(a, b, rtol=RTOL, atol=RTOL * (np.abs(b).max() + 1e-30),
                                   err_msg="call %d differs from dglb200_spmm_csr" % call)
        assert np.array_equal(a, b)                       # same kernels, same slice order
    assert _lib.launch_count() > launches0


def test_static_scratch_rejects_sizes_beyond_the_maximum_and_handles_no_long_rows():
    from dgl_b200 import _lib, ops
    dev = device()
    n, chunk = 400, 32
    src, dst = random_graph(n, 6000, 9, hubs=3, hub_deg=1200, zero_deg_frac=0.1)
    csr = ops.csr_build(torch.from_numpy(dst).to(dev), torch.from_numpy(src).to(dev), n, n)
    plan = csr.plan(chunk)
    feat = SCRATCH_BYTES // 4 // plan.n_chunks + 8        # just beyond the scratch
    feat -= feat % 4
    assert plan.n_chunks * feat * 4 > SCRATCH_BYTES
    st, _, _ = _plan_struct(csr, chunk, feat, with_partial=False)
    x = torch.zeros(n, feat, device=dev)
    y = torch.zeros(n, feat, device=dev)
    assert _call("dglb200_spmm_csr_static_scratch", csr, ops.MSG_COPY_SRC, ops.RED_SUM, x, None, 1, y, st) == -1
    msg = _lib.last_error()
    assert "DGLB200_STATIC_SCRATCH_BYTES" in msg and str(SCRATCH_BYTES) in msg
    # a plan without long rows (or no plan) uses no scratch and simply runs
    st.n_long_rows, st.n_chunks = 0, 0
    x = torch.ones(n, 8, device=dev)
    y = torch.empty(n, 8, device=dev)
    assert _call("dglb200_spmm_csr_static_scratch", csr, ops.MSG_COPY_SRC, ops.RED_SUM, x, None, 1, y, st) == 0
    deg = (csr.indptr[1:] - csr.indptr[:-1]).float()
    assert torch.equal(y[:, 0], deg)
