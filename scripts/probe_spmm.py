"""GPU probe: time the headline SpMM (C2 Reddit-shaped, F=256 fp32) across kernel shapes.
Scratch tool for tuning sweeps; bench.py is the number of record."""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "dgl-1_b200"))

from benchmarks import graphs  # noqa: E402
from dgl_b200 import ops, _lib  # noqa: E402


def timeit(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a = torch.cuda.Event(enable_timing=True)
        b = torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return min(ts), sorted(ts)[len(ts) // 2]


def main():
    scale = float(os.environ.get("PROBE_SCALE", "1.0"))
    feat = int(os.environ.get("PROBE_FEAT", "256"))
    cfg = graphs.C2_REDDIT
    n, e = cfg["n"], int(cfg["e"] * scale)
    dev = torch.device("cuda")
    t0 = time.time()
    src, dst = graphs.powerlaw_symmetric(n, e, cfg["seed"], dev)
    torch.cuda.synchronize()
    print("gen %.2fs" % (time.time() - t0), flush=True)
    t0 = time.time()
    csr = ops.csr_build(dst, src, n, n)
    torch.cuda.synchronize()
    print("csr_build %.3fs" % (time.time() - t0), flush=True)
    deg = ops.degrees(csr)
    print("deg max %d mean %.1f  rows>512: %d  rows>2048: %d" % (
        int(deg.max()), float(deg.float().mean()), int((deg > 512).sum()), int((deg > 2048).sum())))
    del src, dst
    x = torch.randn(n, feat, device=dev)
    lib = _lib.get_lib()
    alg = graphs.spmm_alg_bytes(e, n, feat)
    res = []

    def tune(**kw):
        for k in ("spmm_ucap", "spmm_slab_chunks", "spmm_threads"):
            _lib.check_call(lib.dglb200_set_tuning(k.encode(), int(kw.get(k, 0))))

    ref = None
    configs = [dict()]
    for chunk in (512,):
        for sc in (64, 32, 16):
            for u in (1, 2, 4, 8):
                for nt in (128, 256, 512):
                    configs.append(dict(spmm_ucap=u, spmm_slab_chunks=sc, spmm_threads=nt, chunk=chunk))
    for chunk in (0, 128, 256, 1024, 2048):
        configs.append(dict(chunk=chunk))
    for c in configs:
        chunk = c.get("chunk", 512)
        tune(**c)
        try:
            y = ops.spmm(csr, ops.MSG_COPY_SRC, ops.RED_SUM, x=x, chunk=chunk)[0]
            if ref is None:
                ref = y.clone()
            err = float((y - ref).abs().max())
            best, med = timeit(lambda: ops.spmm(csr, ops.MSG_COPY_SRC, ops.RED_SUM, x=x, chunk=chunk))
        except Exception as ex:  # noqa: BLE001
            print("FAIL", c, ex, flush=True)
            continue
        r = dict(c, ms_best=round(best, 3), ms_med=round(med, 3), alg_gbs=round(alg / best / 1e6, 1),
                 frac=round(alg / best / 1e6 / 6539.5, 3), maxdiff=err)
        res.append(r)
        print(json.dumps(r), flush=True)
    tune()
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "probe_spmm.json"), "w") as f:
        json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
