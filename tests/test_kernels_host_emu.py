"""-m "not gpu": the CUDA KERNEL SOURCES of dgl-1_b200/csrc executed on the CPU.

tests/native/cuda_emu compiles the unmodified .cu / .cuh files for the host against a small
execution model (one host thread per CUDA thread, real rendezvous for warp shuffles /
__match_any_sync / __syncthreads, blocks run one after another) into a library with the same C ABI
as libdgl_b200.so.  This module points `dgl_b200._lib` at that library, lets `ops` accept CPU
tensors, and then runs the BODIES of the `-m gpu` parity tests (a subset of their parameter grids,
sized for a CPU tier of a few minutes) - so the kernels' logic (indexing, pipelines, long-row
splitting, radix sort, argmax ties, bf16 packing, ...) is checked against the oracle and against
the golden vectors of the unmodified reference in a container WITHOUT a GPU.

What this does not show: anything about timing, memory coalescing or the sm_100a code generation;
and the three inline-PTX helpers of common.cuh run as the plain C they stand for.  The product
never loads this library: it is built from tests/ and lives in the temp directory.
"""
import contextlib
import ctypes
import glob
import os

import numpy as np
import pytest
import torch

from tests import test_gpu_csr_build as t_csr
from tests import test_gpu_graph_api as t_api
from tests import test_gpu_sddmm_bucketing as t_sddmm
from tests import test_gpu_spmm as t_spmm
from tests import test_sendrecv_gpu as t_msg
from tests import test_uopv_edge_ops_gpu as t_edge
from tests import test_wire_convert_gpu as t_wire
from tests import test_zz_no_alloc as t_noalloc
from tests import test_fused_gcn_gpu as t_fused
from tests import test_sampling_gpu as t_samp

pytestmark = pytest.mark.timeout(900)     # a lost rendezvous must fail the test, not hang the tier

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
_FACTORIES = ("zeros", "ones", "empty", "full", "randn", "rand", "randint", "arange", "tensor", "as_tensor",
              "zeros_like", "ones_like", "empty_like", "full_like")


@pytest.fixture(scope="module")
def emu_lib():
    import shutil
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    from dgl_b200 import _lib
    from tests.native.cuda_emu import build
    lib = ctypes.CDLL(build.build())
    for name, (res, args) in _lib.SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    assert lib.dglb200_abi_version() == 1
    return lib


def _cpu_device(dev):
    if dev is None:
        return None
    return torch.device("cpu") if torch.device(dev).type == "cuda" else dev


@pytest.fixture()
def emu(monkeypatch, emu_lib):
    """libdgl_b200_emu.so behind dgl_b200.ops; "cuda" means the host for the duration of the test."""
    from dgl_b200 import _lib, graph as graph_mod, ops
    monkeypatch.setattr(_lib, "_LIB", emu_lib)
    monkeypatch.setattr(ops, "_need_cuda", lambda *a: None)
    monkeypatch.setattr(ops, "_stream", lambda: ctypes.c_void_p(0))
    monkeypatch.setattr(torch.cuda, "device", lambda d: contextlib.nullcontext())
    monkeypatch.setattr(torch.cuda, "synchronize", lambda *a, **k: None)
    monkeypatch.setattr(torch.Tensor, "cuda", lambda self, *a, **k: self)
    monkeypatch.setattr(graph_mod.DGLGraph, "_need_device", lambda self: torch.device("cpu"))
    real_to = graph_mod.DGLGraph.to
    monkeypatch.setattr(graph_mod.DGLGraph, "to", lambda self, dev: real_to(self, _cpu_device(dev)))
    for name in _FACTORIES:
        real = getattr(torch, name)

        def wrapped(*a, _real=real, **k):
            if "device" in k:
                k["device"] = _cpu_device(k["device"])
            return _real(*a, **k)
        monkeypatch.setattr(torch, name, wrapped)
    real_gen = torch.Generator
    monkeypatch.setattr(torch, "Generator", lambda device="cpu": real_gen(device=_cpu_device(device)))
    for mod in (t_msg, t_edge, t_noalloc, t_fused, t_samp):
        monkeypatch.setattr(mod, "device", lambda: torch.device("cpu"))
    launches0 = emu_lib.dglb200_launch_count()
    yield emu_lib
    assert emu_lib.dglb200_launch_count() > launches0, "the test launched no emulated kernel"


# ---- kernel (4): CSR build (stable radix sort, scans) --------------------------------------------

_CSR_CASES = [c for c in t_csr.CASES if c[1] is None or max(c[1][0], c[1][1]) <= 60000]   # the rest is GPU-sized


@pytest.mark.parametrize("name,spec", _CSR_CASES, ids=[c[0] for c in _CSR_CASES])
def test_csr_build_kernels(emu, name, spec):
    for transpose in (False, True):
        t_csr.test_csr_mutable_order(name, spec, transpose)
    if name != t_csr.CASES[0][0]:
        t_csr.test_csr_immutable_order(name, spec, 0)
        t_csr.test_csr_immutable_order(name, spec, 1)


def test_csr_build_empty_and_errors(emu):
    # test_gpu_csr_build.test_csr_empty_and_errors minus its "CPU tensors are refused" line, which
    # is exactly what this fixture switches off
    from dgl_b200 import ops, DGLError
    z = torch.zeros(0, dtype=torch.int64)
    csr = ops.csr_build(z, z, 5, 5)
    assert csr.indptr.tolist() == [0] * 6 and csr.nnz == 0
    bad = torch.tensor([0, 7], dtype=torch.int64)
    with pytest.raises(DGLError):
        ops.csr_build(bad, bad, 5, 5)


# ---- kernels (1)(2): SpMM -----------------------------------------------------------------------

def test_spmm_known_answers(emu):
    t_spmm.test_known_answers_hub_graph()


@pytest.mark.parametrize("feat,chunk", [(1, 0), (5, 32), (64, 512), (260, 32), (602, 0)])
def test_spmm_copy_src_sum(emu, feat, chunk):
    t_spmm.test_copy_src_sum(feat, chunk)


@pytest.mark.parametrize("feat,wshape,chunk", [(8, (), 32), (64, (64,), 512), (64, (8,), 32), (60, (5,), 512),
                                               (256, (8,), 32), (7, (7,), 32)])
def test_spmm_src_mul_edge_sum(emu, feat, wshape, chunk):
    t_spmm.test_src_mul_edge_sum(feat, wshape, chunk)


@pytest.mark.parametrize("feat,reduce_op", [(5, "sum"), (64, "max"), (8, "max")])
def test_spmm_copy_edge(emu, feat, reduce_op):
    t_spmm.test_copy_edge(feat, reduce_op)


@pytest.mark.parametrize("feat,msg", [(6, "copy_src"), (64, "src_mul_edge"), (4, "src_mul_edge")])
def test_spmm_max_with_argmax(emu, feat, msg):
    t_spmm.test_max_with_argmax(feat, msg)


@pytest.mark.parametrize("feat", [64, 264])
def test_spmm_bf16(emu, feat):
    t_spmm.test_bf16(feat)


@pytest.mark.parametrize("feat,w_width", [(16, 0), (16, 1), (64, 8), (5, 5)])
def test_spmm_backward_x_sum(emu, feat, w_width):
    t_spmm.test_backward_x_sum(feat, w_width)


def test_spmm_max_propagates_nan(emu):
    t_spmm.test_max_propagates_nan(4)


@pytest.mark.parametrize("feat,chunk", [(8, 0), (128, 64)])
def test_spmm_accumulate_and_bf16_halo(emu, feat, chunk):
    t_spmm.test_accumulate_and_bf16_halo_operand(feat, chunk)


def test_spmm_backward_x_max(emu):
    t_spmm.test_backward_x_max()


@pytest.mark.parametrize("feat,slab", [(96, 24), (12, 5)])
def test_spmm_strided_column_blocks(emu, feat, slab):
    t_spmm.test_strided_column_blocks(feat, slab)


def test_spmm_minus_inf_ragged_heads_and_slot_order(emu, monkeypatch):
    t_spmm.test_max_with_minus_inf_messages(6)
    t_spmm.test_ragged_width_takes_padded_vector_path(41, torch.float32)
    t_spmm.test_multi_head_weights_max_bf16_gradient(8, 3)
    t_spmm.test_slot_ordered_scalar_weights(torch.float32, monkeypatch)


def test_gather_kernels(emu):
    t_spmm.test_gather_rows()
    t_wire.test_gather_rows_convert_bit_exact(13, True)
    t_wire.test_gather_rows_convert_bit_exact(128, False)


# ---- kernel (3): SDDMM, edge softmax, fused GAT; kernel (4b): bucketing, unique ------------------

@pytest.mark.parametrize("heads,dim,op", [(8, 1, "add"), (8, 8, "dot"), (3, 5, "mul"), (4, 32, "dot"), (1, 256, "add"),
                                          (3, 8, "dot")])
def test_sddmm_kernel(emu, heads, dim, op):
    t_sddmm.test_sddmm(heads, dim, op)


def test_sddmm_dot_warp_kernel_equals_generic(emu):
    t_sddmm.test_sddmm_dot_warp_kernel_equals_generic_bitwise(8, 32, "f32")
    t_sddmm.test_sddmm_dot_warp_kernel_equals_generic_bitwise(1, 256, "bf16")
    t_sddmm.test_sddmm_dot_warp_kernel_equals_generic_bitwise(3, 8, "f32")


def test_edge_softmax_and_gat_kernels(emu):
    t_sddmm.test_edge_softmax_fwd_bwd(3)
    t_sddmm.test_gat_aggregate(2, 5, "example")
    t_sddmm.test_gat_aggregate(8, 8, "stable")
    t_sddmm.test_gat_aggregate_head_slabs_bitwise(8, 8, 3, "stable")
    t_sddmm.test_gat_aggregate_head_slabs_bitwise(5, 4, 1, "example")


def test_bucketing_and_unique_kernels(emu):
    for n, m, seed, extra in t_sddmm.BUCKET_CASES:
        if m <= 20000:
            t_sddmm.test_degree_bucketing(n, m, seed, extra)
    t_sddmm.test_degree_bucketing_hub_graph_shapes()
    t_sddmm.test_unique_sorted(10, 30, 0)
    t_sddmm.test_unique_sorted(5, 1, 2)
    t_sddmm.test_unique_sorted(70000, 100, 3)


# ---- the whole stack: host layer + emulated kernels against the reference's golden vectors --------

def test_api_on_emulated_kernels_against_reference_golden(emu, monkeypatch):
    monkeypatch.setattr(t_api, "cu", lambda a: torch.from_numpy(np.ascontiguousarray(a)))
    case = dict(np.load(sorted(glob.glob(os.path.join(GOLDEN, "case_*.npz")))[0]))
    t_api.test_adjacency_and_degrees_bit_exact(case)
    t_api.test_update_all_every_builtin_pair(case)
    t_api.test_max_bit_exact(case)
    t_api.test_gradients(case)
    t_api.test_send_and_recv_pull_push(case)
    t_api.test_gat_example_path(case)
    t_api.test_degree_bucketing_vs_reference(case)


def test_sendrecv_traversal_readout_on_emulated_kernels(emu, monkeypatch):
    monkeypatch.setattr(t_msg, "cu", lambda a: torch.from_numpy(np.ascontiguousarray(a)))
    case = dict(np.load(os.path.join(GOLDEN, "msgpass_0.npz")))
    t_msg.test_send_then_recv_builtin_reducers(case)
    t_msg.test_pair_specifications_on_multigraph(case)
    trav = dict(np.load(os.path.join(GOLDEN, "traversal_0.npz")))
    t_msg.test_frontier_generators_bit_exact(trav)
    bat = dict(np.load(os.path.join(GOLDEN, "batched_0.npz")))
    t_msg.test_batch_readout_unbatch_against_reference(bat)
    t_msg.test_readout_gradient_against_reference(bat)


def test_edge_ops_on_emulated_kernels(emu):
    for op in ("u_add_v", "u_mul_v", "u_dot_v"):
        t_edge.test_sddmm_forward_and_gradients(op, (4, 8))
    t_edge.test_edge_softmax_forward_and_gradient()
    t_edge.test_gat_attention_matches_udf_formulation_with_gradients()
    t_edge.test_gat_attention_slot_order_equals_edge_order()


def test_static_scratch_entry_point_on_emulated_kernels(emu):
    """dglb200_spmm_csr_static_scratch (the operators' allocation-free variant) on the host emulation: the
    `static __device__` scratch is then a host array and cudaGetSymbolAddress its address."""
    t_noalloc.test_static_scratch_spmm_equals_existing_entry_point_twice_in_a_row()
    t_noalloc.test_static_scratch_rejects_sizes_beyond_the_maximum_and_handles_no_long_rows()


def test_fused_gcn_layer_against_reference_example(emu):
    t_fused.test_fused_layer_matches_reference_example_layer(t_fused.GOLDEN[0])
    t_fused.test_epilogue_on_long_rows_ragged_width_and_empty_rows()


# ---- neighbour sampling ------------------------------------------------------------

def test_sampling_golden_cases_of_the_reference(emu):
    t_samp.test_golden_cases_of_the_reference()


@pytest.mark.parametrize("deg,fanout,branch", [(40, 5, "kept set"), (12, 11, "dropped set"), (300, 172, "dropped set"),
                                               (700, 300, "selection")])
def test_sampling_distribution(emu, monkeypatch, deg, fanout, branch):
    monkeypatch.setattr(t_samp, "TRIALS", 240)
    t_samp.test_every_slot_and_every_pair_equally_likely(deg, fanout, branch)


def test_sampling_reproducible_and_hubs(emu, monkeypatch):
    monkeypatch.setattr(t_samp, "HUB_GRAPH", dict(n=700, e=7000, hub_deg=500, fanouts=(
        (10, 2), (100, 1), (450, 1), (1, 3), (0, 2), (5, 0), (10 ** 6, 2))))
    t_samp.test_same_seed_same_sample_and_seeds_differ()
    t_samp.test_hubs_multigraph_and_zero_fanout()
    t_samp.test_bad_seed_is_an_error_and_the_state_recovers()
    t_samp.test_c_abi_too_small_workspace_reports_needed_bytes_and_wipes_the_state()


def test_sampling_reference_sampler_test_with_features(emu):
    # the 1-seed-per-batch loops of the reference's test_sampler.py (300 subgraphs) are left to the GPU tier
    t_samp.test_reference_test_10neighbor_sampler_all_and_features()
