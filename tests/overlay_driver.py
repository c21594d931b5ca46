"""Run as a SUBPROCESS by tests/test_overlay_reference.py (the overlay patches the imported reference
process-wide).  Build container only: needs /root/reference/python and oracle/_ref/libdgl.so.

  python tests/overlay_driver.py emu|gpu

The UNMODIFIED reference (its Python on its compiled C++) computes a set of builtin message-passing calls;
then dgl_b200.overlay.install() patches the reference's four call sites (F.spmm / sparse_matrix*, the
SPMV_WITH_DATA executor, the PackedFunc globals) and the SAME user code runs again - now through the C ABI
into the kernels (emu: the CUDA sources on the host emulation; gpu: libdgl_b200.so) - and must give the
same results and gradients.  Finally the reference's own tests/compute/test_specialization.py and
test_function.py are replayed under the overlay.
"""
import contextlib
import ctypes
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
os.environ.setdefault("DGL_LIBRARY_PATH", os.path.join(ROOT, "oracle", "_ref"))
os.environ.setdefault("DGLBACKEND", "pytorch")
for p in (ROOT, os.path.join(ROOT, "dgl-1_b200"), os.path.join(REF, "python"), os.path.join(REF, "tests"),
          os.path.join(REF, "tests", "compute")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402
import torch  # noqa: E402


# OVERLAY_QUICK=1 skips the six slowest of the reference's tests on the emulated kernels (every index query builds
# its device index on one host thread per CUDA thread); by default all 14 + the 5 sampler tests are replayed
EMU_SKIP = ("test_pull_multi_fallback", "test_spmv_3d_feat", "test_v2v_update_all", "test_e2v_update_all_multi_fn",
            "test_v2v_snr", "test_e2v_recv_multi_fn")


def setup_engine(mode):
    from dgl_b200 import _lib, ops, overlay
    if mode == "emu":
        from tests.native.cuda_emu import build
        lib = ctypes.CDLL(build.build())
        for name, (res, args) in _lib.SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib._LIB = lib
        ops._need_cuda = lambda *a: None
        ops._stream = lambda: ctypes.c_void_p(0)
        torch.cuda.device = lambda d: contextlib.nullcontext()
        overlay.ON_ENGINE = lambda t: True
        overlay.ON_ENGINE_CTX = lambda ctx: True
        return torch.device("cpu")
    return torch.device("cuda")


def cases(dgl, fn, dev):
    """name -> (value, gradient or None) for a set of builtin calls on two graphs."""
    out = {}
    rng = np.random.RandomState(3)
    n, e = 60, 500
    src = rng.randint(0, n, e)
    dst = rng.randint(0, n - 8, e)
    dst[rng.choice(e, 80, replace=False)] = 3                  # hub + zero in-degree tail + parallel edges
    graphs = {"multi": (n, src, dst)}
    hs, hd = [], []
    for i in range(1, 9):
        hs += [0, i]
        hd += [i, 9]
    graphs["hub"] = (10, np.array(hs + [9]), np.array(hd + [0]))
    for gname, (n, src, dst) in graphs.items():
        e = len(src)
        g = dgl.DGLGraph()
        g.add_nodes(n)
        g.add_edges(torch.from_numpy(src), torch.from_numpy(dst))
        gen = torch.Generator().manual_seed(7)
        feats = {"f1": torch.randn(n, generator=gen), "f2": torch.randn(n, 5, generator=gen),
                 "f3": torch.randn(n, 2, 3, generator=gen)}
        w = torch.randn(e, generator=gen)
        for fld, x in feats.items():
            for wname, wt in (("none", None), ("e1", w), ("e2", w.unsqueeze(1))):
                if wname != "none" and fld == "f3":
                    continue
                xd = x.clone().to(dev).requires_grad_(True)
                g.ndata["h"] = xd
                if wt is not None:
                    wd = wt.clone().to(dev).requires_grad_(True)
                    g.edata["w"] = wd
                    mf = fn.src_mul_edge("h", "w", "m")
                else:
                    wd, mf = None, fn.copy_src("h", "m")
                key = "%s/%s/%s" % (gname, fld, wname)
                g.update_all(mf, fn.sum("m", "o"))
                y = g.ndata.pop("o")
                y.sum().backward()
                out[key + "/update_all"] = (y.detach().cpu(), xd.grad.cpu(), None if wd is None else wd.grad.cpu())
                g.ndata["h"] = xd.detach()
                if wd is not None:
                    g.edata["w"] = wd.detach()
                u = torch.from_numpy(src[::3].copy())
                v = torch.from_numpy(dst[::3].copy())
                g.send_and_recv((u, v), mf, fn.sum("m", "o"))
                out[key + "/snr"] = (g.ndata.pop("o").detach().cpu(), None, None)
                g.pull(torch.tensor([3, 0, n - 1, 5]), mf, fn.sum("m", "o"))
                out[key + "/pull"] = (g.ndata.pop("o").detach().cpu(), None, None)
        # e2v (incidence SpMM) and the bucketing fallback
        g.ndata["h"] = feats["f2"].to(dev)
        g.edata["ew"] = torch.randn(e, 5, generator=gen).to(dev)
        g.update_all(fn.copy_edge("ew", "m"), fn.sum("m", "o"))
        out[gname + "/copy_edge_sum"] = (g.ndata.pop("o").cpu(), None, None)
        g.update_all(fn.copy_src("h", "m"), fn.max("m", "o"))
        out[gname + "/copy_src_max"] = (g.ndata.pop("o").cpu(), None, None)
        out[gname + "/in_degrees"] = (g.in_degrees(torch.arange(n)).cpu(), None, None)
    return out


def main():
    mode = sys.argv[1]
    import dgl
    import dgl.function as fn
    from dgl_b200 import _lib, overlay
    dev = setup_engine(mode)
    want = cases(dgl, fn, torch.device("cpu"))               # the unmodified reference
    l0 = _lib.launch_count()
    ov = overlay.install(dgl)
    assert ov.registered >= 31, ov.registered
    got = cases(dgl, fn, dev)
    launches = _lib.launch_count() - l0
    assert launches > 100, "the overlay launched only %d kernels" % launches
    for k in want:
        for a, b in zip(want[k], got[k]):
            if a is None:
                assert b is None
                continue
            a, b = a.double(), b.double()
            tol = 1e-5 * (float(a.abs().max()) + 1e-30)
            assert a.shape == b.shape and float((a - b).abs().max()) <= tol, (k, float((a - b).abs().max()))
    print("OVERLAY_CASES_OK %d cases, %d kernel launches" % (len(want), launches))
    # the reference's own tests, replayed under the overlay (builtin == UDF equivalence, known answers)
    import importlib
    ran = 0
    for modname in ("test_specialization", "test_function"):
        mod = importlib.import_module(modname)
        for name in sorted(dir(mod)):
            if name.startswith("test_") and not (name in EMU_SKIP and os.environ.get("OVERLAY_QUICK")):
                t0 = time.time()
                getattr(mod, name)()
                ran += 1
                if os.environ.get("OVERLAY_TIMES"):
                    print("  %s.%s %.1fs" % (modname, name, time.time() - t0))
    print("REFERENCE_TESTS_OK %d" % ran)
    # the reference's NeighborSampler (python/dgl/contrib/sampling/sampler.py) now draws on the engine: its own
    # tests/compute/test_sampler.py, unchanged
    l1 = _lib.launch_count()
    mod = importlib.import_module("test_sampler")
    names = ["test_10neighbor_sampler_all"] if os.environ.get("OVERLAY_QUICK") else \
        [nm for nm in sorted(dir(mod)) if nm.startswith("test_")]
    for name in names:
        getattr(mod, name)()
    assert _lib.launch_count() > l1 + 50, "the reference sampler did not reach the engine"
    print("REFERENCE_SAMPLER_TESTS_OK %d" % len(names))
    ov.uninstall()
    assert dgl.backend.spmm.__module__.startswith("dgl.")
    print("OVERLAY_OK")


if __name__ == "__main__":
    main()
