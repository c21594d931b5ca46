"""-m "not gpu": the C-ABI library loads and exports every symbol include/dgl_b200.h declares
(no compute calls), error conventions, and the host-side mirror (builtins, scheduler plan)."""
import ctypes
import json
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "dgl_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(dglb200_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from dgl_b200 import _lib
    lib = _lib.get_lib()
    syms = _header_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(lib, s), "libdgl_b200.so does not export %s" % s
    # the ctypes table and the header list the same functions
    assert sorted(_lib.SIGNATURES) == syms
    assert lib.dglb200_abi_version() == 1


def test_plain_c_client_compiles_links_and_runs(tmp_path):
    """The boundary is a C ABI: the header is valid C99 (and C++), and a C program that includes it
    links against libdgl_b200.so and runs the argument-validating calls without a GPU."""
    import shutil
    import subprocess
    from dgl_b200 import _lib
    if shutil.which("gcc") is None:
        pytest.skip("no gcc")
    src = os.path.join(ROOT, "tests", "native", "abi_smoke.c")
    libdir = os.path.dirname(_lib.lib_path())
    exe = str(tmp_path / "abi_smoke")
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror",
                           "-I", os.path.join(ROOT, "include"), src, "-L", libdir, "-ldgl_b200",
                           "-Wl,-rpath," + libdir, "-o", exe])
    out = subprocess.check_output([exe]).decode()
    assert "abi ok" in out
    if shutil.which("g++") is not None:      # the same header from C++
        subprocess.check_call(["g++", "-std=c++11", "-fsyntax-only", "-x", "c++", "-Wall", "-Werror",
                               "-I", os.path.join(ROOT, "include"), src])


def test_converting_gather_kernel_source_runs_under_host_shim(tmp_path):
    """The CUDA source of gather_rows_convert_kernel (no warp primitives) executed on the CPU under a
    minimal host shim: bit-exact float32 -> bfloat16 (RNE) and exact widening over vector and scalar
    paths, strides, gathers and out-of-range ids.  Complements the -m gpu test of the same kernel."""
    import shutil
    import subprocess
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    src = open(os.path.join(ROOT, "dgl-1_b200", "csrc", "gather.cu")).read()
    a = src.index("template <bool TO_BF16>")
    kernel = src[a:src.index("}  // namespace dglb200", a)]
    nat = os.path.join(ROOT, "tests", "native")
    cpp = tmp_path / "emu.cpp"
    cpp.write_text(open(os.path.join(nat, "emu_gather_convert_shim.inc")).read() + kernel +
                   open(os.path.join(nat, "emu_gather_convert_main.inc")).read())
    exe = str(tmp_path / "emu")
    subprocess.check_call(["g++", "-O1", "-std=c++17", "-w", str(cpp), "-o", exe])
    out = subprocess.check_output([exe]).decode()
    assert "fails=0" in out, out


def test_error_convention_no_gpu_needed():
    """0 / -1 + thread-local message (cf. DGLGetLastError, src/runtime/c_runtime_api.cc:130-148)."""
    from dgl_b200 import _lib
    from dgl_b200.base import DGLError
    lib = _lib.get_lib()
    nbytes = ctypes.c_size_t(0)
    assert lib.dglb200_csr_build_workspace_bytes(10, 100, ctypes.byref(nbytes)) == 0
    assert nbytes.value > 0
    assert lib.dglb200_set_tuning(b"no_such_knob", 1) == -1
    assert "no_such_knob" in _lib.last_error()
    with pytest.raises(DGLError):
        _lib.check_call(lib.dglb200_set_tuning(b"no_such_knob", 1))
    # argument validation happens before any CUDA call
    assert lib.dglb200_spmm_csr(-1, 0, 0, None, None, None, 0, 0, 0, None, 4, None, 1, None, None,
                                None, None) == -1
    assert "bad sizes" in _lib.last_error()


def test_ops_refuse_cpu_tensors():
    from dgl_b200 import ops
    from dgl_b200.base import DGLError
    with pytest.raises(DGLError):
        ops.csr_build(torch.zeros(3, dtype=torch.int64), torch.zeros(3, dtype=torch.int64), 4, 4)
    import dgl_b200
    g = dgl_b200.DGLGraph()
    g.add_nodes(3)
    g.add_edges([0, 1], [1, 2])
    g.ndata["h"] = torch.ones(3, 2)
    from dgl_b200 import function as fn
    with pytest.raises(DGLError):
        g.update_all(fn.copy_src("h", "m"), fn.sum("m", "h"))


def test_builtin_surface_matches_reference_fields():
    """python/dgl/function/message.py:47-257, reducer.py:31-97."""
    from dgl_b200 import function as fn
    m = fn.src_mul_edge("h", "w", "m")
    assert (m.src_field, m.edge_field, m.out_field, m.name, m.use_edge_feature) == ("h", "w", "m", "src_mul_edge", True)
    c = fn.copy_src("h", "m")
    assert (c.src_field, c.out_field, c.name, c.use_edge_feature) == ("h", "m", "copy_src", False)
    e = fn.copy_edge("w", "m")
    assert (e.edge_field, e.out_field, e.name, e.use_edge_feature) == ("w", "m", "copy_edge", True)
    for r, nm in ((fn.sum("m", "o"), "sum"), (fn.max("m", "o"), "max")):
        assert (r.msg_field, r.out_field, r.name) == ("m", "o", nm)

    class EB(object):  # UDF-equivalent __call__ (message.py:75-91): rank alignment by appending 1s
        src = {"h": torch.arange(12.).reshape(2, 3, 2)}
        data = {"w": torch.tensor([[2., 3., 4.], [1., 1., 1.]])}
    out = m(EB)["m"]
    assert out.shape == (2, 3, 2) and out[0, 1].tolist() == [6., 9.]

    class NB(object):
        mailbox = {"m": torch.tensor([[[1., 5.], [3., 2.]]])}
    assert fn.sum("m", "o")(NB)["o"].tolist() == [[4., 7.]]
    assert fn.max("m", "o")(NB)["o"].tolist() == [[3., 5.]]


@pytest.mark.parametrize("xs,ws,want", [((5, 8), (9,), 1), ((5, 8), (9, 1), 1), ((5, 8), (9, 8), 8),
                                         ((5, 4, 6), (9, 4, 1), 4), ((5, 5, 5), (9, 5, 5), 25),
                                         ((5,), (9, 3), None), ((5, 4, 6), (9, 1, 6), None)])
def test_weight_broadcast_form(xs, ws, want):
    from dgl_b200.function import _weight_width
    assert _weight_width(xs, ws) == want


def _graph():
    import dgl_b200
    g = dgl_b200.DGLGraph()
    g.add_nodes(4)
    g.add_edges([0, 1, 2, 0], [1, 2, 3, 3])
    g.ndata["h"] = torch.zeros(4, 2)
    g.edata["w"] = torch.zeros(4)
    g.execute = False
    return g


def test_plans_equal_reference_pprint():
    """Executor plans are text-identical to the reference's Prog.pprint() (tests/golden/plans.json,
    generated by oracle/gen_golden.py from the unmodified reference) for update_all."""
    from dgl_b200 import function as fn
    plans = json.load(open(os.path.join(ROOT, "tests", "golden", "plans.json")))
    g = _graph()
    apply_fn = lambda nodes: {"h": nodes.data["h"] * 2}  # noqa: E731
    g.update_all(fn.copy_src("h", "m"), fn.sum("m", "h"))
    assert g.last_prog.lines() == plans["update_all_copy_src_sum"]
    g.update_all(fn.src_mul_edge("h", "w", "m"), fn.sum("m", "h"))
    assert g.last_prog.lines() == plans["update_all_src_mul_edge_sum"]
    g.update_all(fn.copy_src("h", "m"), fn.sum("m", "h"), apply_fn)
    assert g.last_prog.lines() == plans["update_all_copy_src_sum_apply"]
    g.update_all([fn.copy_src("h", "m1"), fn.src_mul_edge("h", "w", "m2")],
                 [fn.sum("m1", "h1"), fn.sum("m2", "h2")])
    assert g.last_prog.lines() == plans["update_all_two_pairs"]
    g.update_all(lambda edges: {"m": edges.src["h"]}, fn.sum("m", "h"))
    assert g.last_prog.lines() == plans["update_all_udf_msg_sum"]
    g.apply_nodes(apply_fn)
    ops_ref = [ln.split("(")[0].split()[-1] for ln in plans["apply_nodes"]]
    assert [ln.split("(")[0].split()[-1] for ln in g.last_prog.lines()] == ops_ref


def test_widened_eligibility_is_one_kernel_executor():
    """max / copy_edge / per-head weights are a single SPMV executor here (the reference sends
    them through EDGE_UDF + degree bucketing, runtime/scheduler.py:540-570)."""
    from dgl_b200 import function as fn
    g = _graph()
    g.update_all(fn.copy_src("h", "m"), fn.max("m", "h"))
    names = [ln.split("(")[0].split()[-1] for ln in g.last_prog.lines()]
    assert names == ["READ_COL", "SPMV", "WRITE_COL_", "WRITE_DICT_"]
    g.update_all(fn.copy_edge("w", "m"), fn.sum("m", "z"))
    names = [ln.split("(")[0].split()[-1] for ln in g.last_prog.lines()]
    assert names == ["READ", "SPMV", "WRITE_COL_", "WRITE_DICT_"]
    with pytest.raises(Exception):
        g.update_all(fn.copy_src("h", "m"), fn.sum("nope", "h"))
