"""-m "not gpu": the HOST layer (DGLGraph entry points, scheduler, plans, write-back, autograd
wiring) against the golden vectors of the unmodified reference, with the kernel-launching functions
of dgl_b200.ops swapped for the CPU restatements of tests/cpu_engine.py.

The bodies are the `-m gpu` tests of tests/test_gpu_graph_api.py, run unchanged: only the device of
the tensors and the engine behind `ops` differ.  This checks everything above the C ABI in a
container without a GPU; the kernels themselves are checked by the `-m gpu` suite on the B200.
"""
import glob
import os

import numpy as np
import pytest
import torch

from tests import cpu_engine
from tests import test_gpu_graph_api as api
from tests import test_sendrecv_gpu as mp

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = sorted(glob.glob(os.path.join(GOLDEN, "case_*.npz")))

BODIES = [api.test_adjacency_and_degrees_bit_exact, api.test_update_all_every_builtin_pair,
          api.test_max_bit_exact, api.test_gradients, api.test_send_and_recv_pull_push,
          api.test_udf_fallbacks_equal_builtins, api.test_degree_bucketing_vs_reference]


def _cpu_graph(case, extra_loops=False):
    import dgl_b200
    n, src, dst = int(case["n"]), case["src"], case["dst"]
    if extra_loops:
        src = np.concatenate([src, np.arange(n)])
        dst = np.concatenate([dst, np.arange(n)])
    g = dgl_b200.DGLGraph()
    g.add_nodes(n)
    g.add_edges(torch.from_numpy(src), torch.from_numpy(dst))
    return g.to("cpu")


@pytest.fixture()
def cpu_api(monkeypatch):
    monkeypatch.setattr(api, "cu", lambda a: torch.from_numpy(np.ascontiguousarray(a)))
    monkeypatch.setattr(api, "graph", _cpu_graph)
    with cpu_engine.installed():
        yield api


@pytest.mark.parametrize("body", BODIES, ids=[b.__name__ for b in BODIES])
@pytest.mark.parametrize("path", CASES[:2], ids=[os.path.basename(c) for c in CASES[:2]])
def test_host_layer_against_reference_golden(cpu_api, path, body):
    body(dict(np.load(path)))


MP_CASES = sorted(glob.glob(os.path.join(GOLDEN, "msgpass_*.npz")))
MP_BODIES = [mp.test_send_then_recv_builtin_reducers, mp.test_send_then_recv_udf_reducer_with_apply,
             mp.test_consecutive_sends_partial_recv, mp.test_send_all_recv_all_equals_update_all,
             mp.test_recv_without_pending_messages_is_apply, mp.test_pair_specifications_on_multigraph,
             mp.test_push_with_apply, mp.test_subset_apply_nodes_and_edges, mp.test_user_initializers,
             mp.test_user_facing_sparse_matrices, mp.test_structure_changes_after_drop_coo]


@pytest.fixture()
def cpu_mp(monkeypatch):
    monkeypatch.setattr(mp, "cu", lambda a: torch.from_numpy(np.ascontiguousarray(a)))
    monkeypatch.setattr(mp, "device", lambda: torch.device("cpu"))
    with cpu_engine.installed():
        yield mp


@pytest.mark.parametrize("body", MP_BODIES, ids=[b.__name__ for b in MP_BODIES])
@pytest.mark.parametrize("path", MP_CASES, ids=[os.path.basename(c) for c in MP_CASES])
def test_send_recv_host_layer_against_reference_golden(cpu_mp, path, body):
    body(dict(np.load(path)))


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "traversal_*.npz"))),
                         ids=lambda p: os.path.basename(p))
def test_traversal_and_propagation_against_reference_golden(cpu_mp, path):
    case = dict(np.load(path))
    mp.test_frontier_generators_bit_exact(case)
    mp.test_propagation_against_reference(case)


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "batched_*.npz"))),
                         ids=lambda p: os.path.basename(p))
def test_batched_graph_readout_against_reference_golden(cpu_mp, path):
    case = dict(np.load(path))
    mp.test_batch_readout_unbatch_against_reference(case)
    mp.test_readout_gradient_against_reference(case)


def test_edge_ops_autograd_wiring(monkeypatch):
    from tests import test_uopv_edge_ops_gpu as eo
    monkeypatch.setattr(eo, "device", lambda: torch.device("cpu"))
    with cpu_engine.installed():
        for op in ("u_add_v", "u_mul_v", "u_dot_v", "copy_u", "copy_v"):
            for tail in ((6,), (4, 8), (3, 1)):
                eo.test_sddmm_forward_and_gradients(op, tail)
        eo.test_edge_softmax_forward_and_gradient()
        eo.test_gat_attention_matches_udf_formulation_with_gradients()
        eo.test_apply_edges_builtin_is_one_sddmm_executor()


def test_reference_style_scenarios_on_host_layer(cpu_mp):
    mp.test_recv_0deg_like_reference()
    mp.test_send_multigraph_like_reference()


def test_cpu_engine_is_not_left_installed():
    """Outside the context manager the product has no CPU path."""
    from dgl_b200 import ops
    from dgl_b200.base import DGLError
    with cpu_engine.installed():
        pass
    with pytest.raises(DGLError):
        ops.csr_build(torch.zeros(3, dtype=torch.int64), torch.zeros(3, dtype=torch.int64), 4, 4)
