"""-m gpu: kernel (4) CSR build vs the oracle, bit-exact (integer work).
Mirrors what tests/compute/test_graph_index.py:29-112 checks in the reference (adjacency of
mutable vs immutable index, degrees) at the GetAdj surface."""
import numpy as np
import pytest
import torch

from oracle import oracle
from tests.util import random_graph, hub_graph

pytestmark = pytest.mark.gpu


def _build(src, dst, n, transpose, order):
    from dgl_b200 import ops
    s = torch.from_numpy(src).cuda()
    d = torch.from_numpy(dst).cuda()
    row, col = (s, d) if transpose else (d, s)
    return ops.csr_build(row, col, n, n, order=order)


CASES = [
    ("hub", None),
    ("rand_small", (50, 300, 0)),
    ("rand_multi", (20, 2000, 1)),
    ("rand_zero_deg", (1000, 5000, 2)),
    ("rand_big", (70000, 900000, 3)),
    ("one_node", (1, 40, 4)),
]


@pytest.mark.parametrize("name,spec", CASES)
@pytest.mark.parametrize("transpose", [False, True])
def test_csr_mutable_order(name, spec, transpose):
    from dgl_b200 import ops
    if spec is None:
        src, dst = hub_graph()
        n = 10
    else:
        n, e, seed = spec
        src, dst = random_graph(n, e, seed, hubs=2 if n > 10 else 0, hub_deg=n // 2,
                                zero_deg_frac=0.3 if "zero" in name else 0.0)
    csr = _build(src, dst, n, transpose, 0)
    indptr, indices, eid = oracle.get_adj_csr(src, dst, n, transpose)
    assert np.array_equal(csr.indptr.cpu().numpy(), indptr)
    assert np.array_equal(ops.widen(csr.indices).cpu().numpy(), indices)
    assert np.array_equal(ops.widen(csr.eids).cpu().numpy(), eid)
    deg = ops.degrees(csr).cpu().numpy()
    assert np.array_equal(deg, np.diff(indptr))
    rows = ops.csr_row_ids(csr).cpu().numpy()
    assert np.array_equal(rows, np.repeat(np.arange(n), np.diff(indptr)))


@pytest.mark.parametrize("name,spec", CASES[1:])
@pytest.mark.parametrize("sort_on", [0, 1])
def test_csr_immutable_order(name, spec, sort_on):
    from dgl_b200 import ops
    n, e, seed = spec
    src, dst = random_graph(n, e, seed)
    csr = _build(src, dst, n, sort_on == 0, 1)
    indptr, indices, eid = oracle.immutable_csr(src, dst, n, sort_on)
    assert np.array_equal(csr.indptr.cpu().numpy(), indptr)
    assert np.array_equal(ops.widen(csr.indices).cpu().numpy(), indices)
    assert np.array_equal(ops.widen(csr.eids).cpu().numpy(), eid)


def test_csr_empty_and_errors():
    from dgl_b200 import ops, DGLError
    z = torch.zeros(0, dtype=torch.int64, device="cuda")
    csr = ops.csr_build(z, z, 5, 5)
    assert csr.indptr.cpu().tolist() == [0] * 6 and csr.nnz == 0
    bad = torch.tensor([0, 7], dtype=torch.int64, device="cuda")
    with pytest.raises(DGLError):
        ops.csr_build(bad, bad, 5, 5)
    with pytest.raises(DGLError):
        ops.csr_build(bad.cpu(), bad.cpu(), 9, 9)
