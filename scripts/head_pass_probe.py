"""src_mul_edge -> sum with per-head weights (E, 8) at C2 size: one head per pass with a scalar weight (the default when
a head is a power-of-two number of 16-byte chunks) against the usual L2 slabs with a weight per chunk."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "dgl-1_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)
import torch  # noqa: E402

from benchmarks import graphs  # noqa: E402
from benchmarks.configs_bench import timeit  # noqa: E402
from dgl_b200 import _lib, ops  # noqa: E402

DEV = torch.device("cuda")
c = graphs.C2_REDDIT
n, e = c["n"], c["e"]
src, dst = graphs.powerlaw_symmetric(n, e, c["seed"], DEV)
cin = ops.csr_build(dst, src, n, n)
cslot = ops.CSR(n, n, cin.indptr, cin.indices, None)
del src, dst
lib = _lib.get_lib()
w = torch.rand(e, 8, device=DEV)
for hid in (16, 32):
    x = torch.randn(n, 8 * hid, device=DEV)
    ref = None
    for mode in (1, 2):
        _lib.check_call(lib.dglb200_set_tuning(b"spmm_head_pass", mode))
        ms = timeit(lambda: ops.spmm(cslot, ops.MSG_SRC_MUL_EDGE, ops.RED_SUM, x=x, w=w, w_width=8), warm=2, iters=5)
        y = ops.spmm(cslot, ops.MSG_SRC_MUL_EDGE, ops.RED_SUM, x=x, w=w, w_width=8)[0]
        if ref is None:
            ref = y
        print("F=%d head_pass=%d: %.2f ms, max |diff| vs mode 1 = %.3g" % (8 * hid, mode, ms, float((y - ref).abs().max())))
    ms = timeit(lambda: ops.spmm(cslot, ops.MSG_COPY_SRC, ops.RED_SUM, x=x), warm=2, iters=5)
    print("F=%d copy_src (no weights): %.2f ms" % (8 * hid, ms))
_lib.check_call(lib.dglb200_set_tuning(b"spmm_head_pass", 0))
