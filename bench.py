"""bench.py - the number of record for the hot path.

One "step" = one pass of the hot path over one batch of synthetic input: the GCN aggregation
`update_all(copy_src, sum)` forward (in-edge CSR SpMM) + its gradient (the transposed SpMM
over the out-edge CSR), fp32, on the workload BASELINE.json's metric is quoted on:

  C2  Reddit-shaped graph, N=232,965, E=114,615,892, aggregated width F=256  (default, all N)
  C5  RMAT graph, N=50M, E=150M, F=128 (`--workload c5`; the multi-GPU config of BASELINE.json)

N GPUs (`--gpus N`, launched by torchrun, one rank per GPU): destination rows are 1-D
partitioned into N contiguous edge-balanced blocks; each rank owns the rows of X, Y, dY, dX of
its block; the halo exchange is an NCCL all-gather of the feature blocks, then the local SpMM
runs over the rank's row block.  Strong scaling: the graph is fixed as N grows.

Prints ONE JSON line (rank 0).  `value` is whole-job ALGORITHMIC GB/s of the SpMM aggregation
(gather model of SURVEY.md 8d: E*(F*4+4) + N*F*4 + (N+1)*8 bytes per SpMM, two SpMMs per step)
with inputs resident in HBM; `e2e` is the same metric through the host API (DGLGraph.update_all
+ backward) with HOST (pinned) feature buffers, H2D / D2H copies inside the timed region.

`--impl reference` times the reference's own CPU implementation of the path - the call
`th.sparse.mm(uncoalesced COO adjacency, X)` + its autograd backward, exactly what
SPMVExecutor.run -> F.spmm does (python/dgl/runtime/ir/executor.py:452-473;
python/dgl/backend/pytorch/tensor.py:45-51,143-146) - on the box's host cores, on a bounded
sample of the same workload.  torch is the third-party dependency that holds the arithmetic in
the reference; oracle/ provides the adjacency restatement (Graph::GetAdj "coo").
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "dgl-1_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402

from benchmarks import graphs  # noqa: E402

METRIC = "spmm_aggregation_alg_GBps"
UNIT = "GB/s"


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback"


def workload_cfg(name):
    if name == "c2":
        c = graphs.C2_REDDIT
        return dict(name="C2 reddit-shaped GCN aggregation copy_src->sum fwd+bwd",
                    n=c["n"], e=c["e"], feat=c["hidden"], seed=c["seed"], kind="powerlaw")
    if name == "c5":
        c = graphs.C5_RMAT
        return dict(name="C5 RMAT SSE/GCN aggregation copy_src->sum fwd+bwd",
                    n=c["n"], e=c["e"], feat=c["feat"], seed=c["seed"], kind="rmat")
    if name == "tiny":  # CI / debugging only
        return dict(name="tiny powerlaw", n=20000, e=2000000, feat=256, seed=1, kind="powerlaw")
    raise SystemExit("unknown workload %s" % name)


def make_edges(cfg, device, e=None, scale_deg=1.0):
    e = cfg["e"] if e is None else e
    if cfg["kind"] == "powerlaw":
        return graphs.powerlaw_symmetric(cfg["n"], e, cfg["seed"], device,
                                         max_deg=max(64, int(22000 * scale_deg)))
    return graphs.rmat(26, e, cfg["seed"], device, n=cfg["n"])


# ---------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: torch.sparse.mm on the uncoalesced COO, CPU
# ---------------------------------------------------------------------------------------------

def cpu_reference(cfg, steps, warmup, budget_s=25.0, gcn=False, full=False):
    """Times fwd+bwd of th.sparse.mm on the host cores.

    full=False (the `cpu_baseline` leg of the B200 arm): a bounded sample - same N and F, E/k edges of
    the same generator, k sized so that steps+warmup passes take about `budget_s` seconds.
    full=True (`--impl reference`): the WHOLE edge list of the configuration (same_config), as many
    timed passes (>= 1, <= steps) as fit `budget_s` after one warm-up pass; if even one pass would not
    fit 4 x budget_s the sampled mode is used and says so."""
    from oracle import oracle
    n, feat = cfg["n"], cfg["feat"]
    # torchrun exports OMP_NUM_THREADS=1: pin the thread count so the baseline is the same at every N
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    cores = torch.get_num_threads()
    # probe the machine with a small slice to size the sample
    probe_e = min(cfg["e"], 2_000_000)
    src, dst = make_edges(cfg, "cpu", probe_e, probe_e / cfg["e"])

    busy = [0.0]

    def one_pass(src, dst, reps):
        idx, _ = oracle.get_adj_coo(src.numpy(), dst.numpy())          # Graph::GetAdj "coo"
        e = src.shape[0]
        adj = torch.sparse_coo_tensor(torch.from_numpy(idx).view(2, e), torch.ones(e), (n, n))
        x = torch.randn(n, feat, requires_grad=True)
        dy = torch.randn(n, feat)
        ts = []
        c0 = time.process_time()
        for _ in range(reps):
            t0 = time.perf_counter()
            y = torch.sparse.mm(adj, x)
            y.backward(dy)
            ts.append(time.perf_counter() - t0)
            x.grad = None
        busy[0] = (time.process_time() - c0) / max(sum(ts), 1e-9)     # CPU seconds per wall second
        return ts

    t_probe = min(one_pass(src, dst, 2))
    per_edge = t_probe / probe_e
    mode = "sample"
    if full and per_edge * cfg["e"] <= 4 * budget_s:
        mode = "full"
        e_s = cfg["e"]
        timed = int(max(1, min(steps, budget_s / (per_edge * e_s) - 1)))
        warm = 1
        src, dst = make_edges(cfg, "cpu")
        ts = one_pass(src, dst, warm + timed)[warm:]
    else:
        total = max(1, steps + warmup)
        e_s = int(min(cfg["e"], max(probe_e, budget_s / total / per_edge)))
        e_s -= e_s % 2
        src, dst = make_edges(cfg, "cpu", e_s, e_s / cfg["e"])
        ts = one_pass(src, dst, total)[warmup:] or [t_probe]
    t = sum(ts) / len(ts)
    alg = 2 * graphs.spmm_alg_bytes(e_s, n, feat)
    gcn = (gcn_epoch_cpu(src, dst, n, cfg, epochs=1 if mode == "full" else 2, warm=0 if mode == "full" else 1)
           if (gcn and cfg["kind"] == "powerlaw") else None)
    return dict(gcn_epoch=gcn, value=alg / t / 1e9, unit=UNIT, cores=cores, kind="port",
                sample="th.sparse.mm(COO)+autograd bwd, fp32, N=%d F=%d, E=%d of %d edges (%s), "
                       "%d timed passes, %.3f s/pass; %d torch threads offered, %.1f busy on average"
                       % (n, feat, e_s, cfg["e"], "the full configuration" if e_s == cfg["e"]
                          else "1/%.1f sample" % (cfg["e"] / e_s), len(ts), t, cores, busy[0]),
                effective_cores=round(busy[0], 2), same_config=bool(e_s == cfg["e"]), timed_passes=len(ts),
                edge_feat_per_s=2 * e_s * feat / t, ms_per_step=t * 1e3, sample_edges=e_s)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg = workload_cfg(args.workload)
    r = cpu_reference(cfg, args.steps, args.warmup, budget_s=args.ref_budget, gcn=not args.no_gcn,
                      full=not args.ref_sample)
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": cfg["name"], "n": cfg["n"], "e": cfg["e"], "feat": cfg["feat"],
                   "sample_edges": r["sample_edges"], "same_config": r["same_config"],
                   "timed_passes": r["timed_passes"],
                   "note": "every pass is the whole step (fwd+bwd over the edge list named in sample_edges); "
                           "the number of timed passes is bounded by --ref-budget seconds, not by --steps"},
        "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "effective_cores", "kind", "sample")},
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "edge_feat_per_s": r["edge_feat_per_s"], "gpu_launches": 0, "gcn_epoch": r["gcn_epoch"],
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------
# GCN epoch (BASELINE.json metric, second half): the reference's examples/pytorch/gcn/gcn_spmv.py
# model - per layer dropout, h @ W, * norm, update_all(copy_src, sum), * norm, + bias, activation -
# two layers feat_in -> hidden -> classes, cross-entropy, Adam; one epoch = forward + backward +
# optimizer step over the full graph.  Dense GEMMs are torch (cuBLAS / MKL) in both arms; the
# aggregation is this engine's update_all in the B200 arm and th.sparse.mm in the reference arm.
# ---------------------------------------------------------------------------------------------

class GCNModel(object):
    def __init__(self, feat_in, hidden, classes, device, dropout=0.5):
        gen = torch.Generator().manual_seed(7)          # same weights on every rank
        def uni(*shape):
            stdv = 1.0 / (shape[-1] ** 0.5)             # reset_parameters of the example (gcn_spmv.py:35-39)
            return ((torch.rand(*shape, generator=gen) * 2 - 1) * stdv).to(device).requires_grad_(True)
        self.w1, self.b1 = uni(feat_in, hidden), uni(hidden)
        self.w2, self.b2 = uni(hidden, classes), uni(classes)
        self.params = [self.w1, self.b1, self.w2, self.b2]
        self.dropout = dropout
        self.opt = torch.optim.Adam(self.params, lr=1e-2, weight_decay=5e-4)

    def forward(self, x, norm, aggregate, fused_layer=None):
        drop = torch.nn.functional.dropout
        if fused_layer is not None:
            # dgl_b200.fused: destination norm, bias and ReLU in the SpMM's store epilogue, source norm in the
            # transposed SpMM's epilogue on the way back (examples/pytorch/gcn/gcn_spmv.py:45-62 in 2 kernels)
            h = fused_layer(torch.mm(drop(x, self.dropout), self.w1), self.b1, True)
            return fused_layer(torch.mm(drop(h, self.dropout), self.w2), self.b2, False)
        h = torch.mm(drop(x, self.dropout), self.w1) * norm
        h = torch.relu(aggregate(h, "h") * norm + self.b1)
        h = torch.mm(drop(h, self.dropout), self.w2) * norm
        return aggregate(h, "c") * norm + self.b2

    def epoch(self, x, labels, norm, aggregate, n_total, allreduce=None, fused_layer=None):
        logits = self.forward(x, norm, aggregate, fused_layer)
        loss = torch.nn.functional.cross_entropy(logits, labels, reduction="sum") / n_total
        self.opt.zero_grad(set_to_none=True)
        loss.backward()
        if allreduce is not None:                       # data-parallel: sum the per-rank weight gradients
            flat = torch.cat([p.grad.reshape(-1) for p in self.params])
            allreduce(flat)
            o = 0
            for p in self.params:
                p.grad.copy_(flat[o:o + p.numel()].view_as(p))
                o += p.numel()
        self.opt.step()
        return loss


def gcn_dims(cfg):
    c = graphs.C2_REDDIT
    return c["feat_in"], cfg["feat"], c["classes"]


def gcn_epoch_cpu(src, dst, n, cfg, epochs=2, warm=1):
    """Reference arm: the same model with th.sparse.mm on the uncoalesced COO (CPU)."""
    from oracle import oracle
    feat_in, hidden, classes = gcn_dims(cfg)
    e = src.shape[0]
    idx, _ = oracle.get_adj_coo(src.numpy(), dst.numpy())
    adj = torch.sparse_coo_tensor(torch.from_numpy(idx).view(2, e), torch.ones(e), (n, n))
    deg = torch.bincount(dst, minlength=n).float()
    norm = deg.pow(-0.5)
    norm[torch.isinf(norm)] = 0
    norm = norm.unsqueeze(1)
    x = torch.randn(n, feat_in)
    labels = torch.randint(0, classes, (n,))
    model = GCNModel(feat_in, hidden, classes, "cpu")
    agg = lambda h, tag: torch.sparse.mm(adj, h)  # noqa: E731
    ts = []
    for i in range(warm + epochs):
        t0 = time.perf_counter()
        model.epoch(x, labels, norm, agg, n)
        ts.append(time.perf_counter() - t0)
    ts = ts[warm:]
    t = sum(ts) / len(ts)
    return {"ms": t * 1e3, "edges": int(e), "edges_per_s": e / t, "epochs": len(ts),
            "model": "GCN %d-%d-%d, dropout 0.5, Adam" % (feat_in, hidden, classes)}


# ---------------------------------------------------------------------------------------------
# clocks
# ---------------------------------------------------------------------------------------------

class ClockSampler(object):
    """nvidia-smi polled every 20 ms into a file.  Started BEFORE the warm-up (the tool needs a few
    hundred ms to come up), `mark()` is called when the timed region starts; `stop()` reports the
    samples taken after the mark, or - when the timed region was shorter than a polling interval -
    the last samples of the warm-up, which ran the same kernels back to back (`window` says which)."""
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")
    NAMES = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        self.mark_line = 0
        try:
            self.p = subprocess.Popen(
                ["nvidia-smi", "--query-gpu=" + self.FIELDS, "--format=csv,noheader,nounits",
                 "-lms", "20", "-i", str(index)], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def _lines(self):
        try:
            with open(self.f.name) as f:
                return f.read().splitlines()
        except OSError:
            return []

    def mark(self):
        self.mark_line = len(self._lines())

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        lines = self._lines()
        self.f.close()
        os.unlink(self.f.name)
        window = "timed"
        picked = lines[self.mark_line:]
        if len(picked) < 2:                      # timed region shorter than the polling interval
            window = "warmup_tail+timed"
            picked = lines[max(0, self.mark_line - 10):]
        sm, mx, reasons, power = [], [], set(), []
        for ln in picked:
            parts = [s.strip() for s in ln.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
                power.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(self.NAMES, parts[3:7]):
                if val == "Active":
                    reasons.add(name)
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(mx), reasons=sorted(reasons),
                       samples=len(sm), power_w_max=max(power), window=window)
        return out


# ---------------------------------------------------------------------------------------------
# B200 arm
# ---------------------------------------------------------------------------------------------

class numa_local(object):
    """Allocate the pinned host buffers of the e2e legs on the NUMA node of the GPU: while inside, the process runs
    on the GPU's own CPUs (NVML's affinity mask), so first-touch places the pages next to the PCIe root the GPU hangs
    off; the previous affinity comes back afterwards (the CPU baselines keep all cores).  The pool's boxes are
    16-vCPU guests that expose ONE affinity domain, so there it changes nothing ("single affinity domain") - the
    host<->device rate still differs between boxes (e2e 13.9 - 14.1 ms per C2 step on most, 32.4 ms on one, same
    code and clocks: placement below the guest, out of reach).  Best effort; `.note` says what was done."""

    def __init__(self, dev):
        self.dev, self.old, self.note = dev, None, "default placement"

    def __enter__(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            idx = self.dev.index if self.dev.index is not None else torch.cuda.current_device()
            props = torch.cuda.get_device_properties(idx)
            h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + str(props.uuid)).encode()) if hasattr(props, "uuid") \
                else pynvml.nvmlDeviceGetHandleByIndex(idx)
            words = (os.cpu_count() + 63) // 64
            mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
            cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (int(m) >> b) & 1}
            allowed = os.sched_getaffinity(0)
            cpus &= allowed
            if cpus and cpus != allowed:
                self.old = allowed
                os.sched_setaffinity(0, cpus)
                self.note = "numa-local to the GPU (%d of %d cpus during allocation)" % (len(cpus), len(allowed))
            elif cpus:
                self.note = "single affinity domain (%d cpus)" % len(cpus)
        except Exception as ex:  # noqa: BLE001
            self.note = "default placement (%s)" % str(ex)[:80]
        return self

    def __exit__(self, *exc):
        if self.old is not None:
            try:
                os.sched_setaffinity(0, self.old)
            except Exception:  # noqa: BLE001
                pass
        return False


def e2e_partitioned(part, x_loc, dy_loc, y_loc, dx_loc, alg_step, barrier, dev, world, args):
    """N > 1 end-to-end: every rank's row blocks of X and dY start in HOST (pinned) memory and its
    blocks of Y and dX end there, every step, through the partition's public calls
    (RowPartition / HaloPartition .forward / .backward = exchange + SpMM).  Same three-stream
    pipeline as the N = 1 e2e: H2D on one side stream, D2H on another (PCIe is full duplex), at most
    two steps in flight; the device copies that refill the exchange buffers and the kernels are
    stream-ordered on the current stream, and the exchange's own signalling keeps a block from being
    overwritten while a peer still reads it.  Timed with CUDA events, max over ranks."""
    import torch.distributed as dist
    rows, feat = x_loc.shape
    with numa_local(dev) as placement:
        xh = torch.randn(rows, feat).pin_memory()
        dyh = torch.randn(rows, feat).pin_memory()
        yh = torch.empty(rows, feat).pin_memory()
        dxh = torch.empty(rows, feat).pin_memory()
    s_in, s_out = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
    cur = torch.cuda.current_stream(dev)
    stage = [(torch.empty_like(x_loc), torch.empty_like(dy_loc), torch.empty_like(y_loc), torch.empty_like(dx_loc))
             for _ in range(2)]
    done = []

    def one(i):
        xs, dys, ys, dxs = stage[i % 2]
        with torch.cuda.stream(s_in):
            xs.copy_(xh, non_blocking=True)
            ev_x = torch.cuda.Event()
            ev_x.record(s_in)
            dys.copy_(dyh, non_blocking=True)
            ev_dy = torch.cuda.Event()
            ev_dy.record(s_in)
        cur.wait_event(ev_x)
        x_loc.copy_(xs, non_blocking=True)
        part.forward(x_loc, out=ys)
        ev_y = torch.cuda.Event()
        ev_y.record(cur)
        with torch.cuda.stream(s_out):
            s_out.wait_event(ev_y)
            yh.copy_(ys, non_blocking=True)
        cur.wait_event(ev_dy)
        dy_loc.copy_(dys, non_blocking=True)
        part.backward(dy_loc, out=dxs)
        ev_dx = torch.cuda.Event()
        ev_dx.record(cur)
        with torch.cuda.stream(s_out):
            s_out.wait_event(ev_dx)
            dxh.copy_(dxs, non_blocking=True)
            d = torch.cuda.Event()
            d.record(s_out)
        done.append(d)
        if len(done) >= 2:                   # the staging pair of step i-1 is reused by step i+1
            done.pop(0).synchronize()

    for i in range(2):
        one(i)
    barrier()
    done.clear()
    k = max(3, min(args.steps, 10))
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(k):
        one(i)
    torch.cuda.current_stream(dev).wait_stream(s_out)
    b.record()
    barrier()
    t = torch.tensor([a.elapsed_time(b) / k], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    nbytes = torch.tensor([2.0 * rows * feat * 4], device=dev, dtype=torch.float64)
    dist.all_reduce(nbytes)                                   # whole job, all ranks
    return {"value": alg_step / (ms * 1e-3) / 1e9, "unit": UNIT,
            "h2d_bytes_per_step": int(nbytes.item()), "d2h_bytes_per_step": int(nbytes.item()),
            "ms_per_step": ms, "steps": k, "method": "3-stream pipeline, 2 steps in flight (same as N=1)",
            "host_buffers": placement.note}


# ---------------------------------------------------------------------------------------------
# self-checks: every driver-run line verifies the configuration it times (outside the timed region)
# ---------------------------------------------------------------------------------------------

def _sample_rows(csr, k, gen_seed=5):
    """k row ids of a CSR: the highest-degree rows (hubs: the long-row slices + fix-up path), a few
    empty rows, and random ones."""
    deg = csr.indptr[1:] - csr.indptr[:-1]
    n = int(deg.shape[0])
    k = min(k, n)
    hubs = torch.topk(deg, min(32, k))[1]
    g = torch.Generator(device=deg.device)
    g.manual_seed(gen_seed)
    rnd = torch.randint(0, n, (k - hubs.numel(),), device=deg.device, generator=g)
    return torch.unique(torch.cat([hubs, rnd]))


def _rows_coo(csr, rows):
    """(col ids, local row id) of every slot of the given rows, in slot (= edge id) order."""
    lens = (csr.indptr[rows + 1] - csr.indptr[rows])
    first = csr.indptr[rows]
    local = torch.repeat_interleave(torch.arange(rows.numel(), device=rows.device), lens)
    start = torch.cumsum(lens, 0) - lens
    slot = first[local] + (torch.arange(int(lens.sum()), device=rows.device) - start[local])
    return csr.indices[slot].long(), local


def check_rows_fp64(csr, x, y_rows_of, rows, rtol=1e-5):
    """Sampled rows of y = A x against an independent computation: torch gather + index_add in
    float64 on the device (no kernel of this library).  y_rows_of(rows) returns the rows under test."""
    cols, local = _rows_coo(csr, rows)
    want = torch.zeros((rows.numel(), x.shape[1]), dtype=torch.float64, device=x.device)
    for s0 in range(0, cols.numel(), 1 << 22):
        want.index_add_(0, local[s0:s0 + (1 << 22)], x[cols[s0:s0 + (1 << 22)]].double())
    got = y_rows_of(rows).double()
    scale = float(want.abs().max()) + 1e-30
    return bool(((got - want).abs() <= rtol * scale + rtol * want.abs()).all())


def check_rows_oracle(csr, x_cpu, y_rows, rows, rtol=1e-5):
    """Sampled rows against oracle/c_oracle.spmm_sum (the plain-C restatement, summation in edge-id
    order like th.sparse.mm on the reference's COO)."""
    from oracle import c_oracle
    if not c_oracle.available():
        return None
    import numpy as np
    cols, local = _rows_coo(csr, rows)
    want = c_oracle.spmm_sum(cols.cpu().numpy(), local.cpu().numpy(), int(rows.numel()), "copy_src", x=x_cpu)
    got = y_rows.cpu().numpy()
    scale = float(np.abs(want).max()) + 1e-30
    return bool((np.abs(got - want) <= rtol * scale + rtol * np.abs(want)).all())


def check_single_gpu(loc_in, loc_out, x, dy, y, dx, feat, with_oracle):
    """N = 1: (a) the result of the configuration that was timed (at C2: F = 256 in four L2 slabs along
    blockIdx.y) bit for bit against column-block calls that take the un-slabbed path
    (dglb200_spmm_csr_strided over 64-float blocks); (b) sampled rows (hubs included) of Y and dX
    against the C oracle (when the operand fits the host) and against a float64 torch computation."""
    from dgl_b200 import ops
    res = {}
    blk = 64
    if feat > blk and feat % blk == 0 and x.shape[0] * feat * 4 <= (4 << 30):
        ok = True
        for csr_, src_, out_ in ((loc_in, x, y), (loc_out, dy, dx)):
            ref = torch.empty_like(out_)
            for s0 in range(0, feat, blk):
                ops.spmm(csr_, ops.MSG_COPY_SRC, ops.RED_SUM, x=src_[:, s0:s0 + blk], out=ref[:, s0:s0 + blk])
            ok = ok and torch.equal(ref, out_)
            del ref
        res["slab_path_equals_column_blocks_bitwise"] = bool(ok)
    rows_in, rows_out = _sample_rows(loc_in, 2000), _sample_rows(loc_out, 2000)
    res["sampled_rows_fp64"] = bool(check_rows_fp64(loc_in, x, lambda r: y[r], rows_in) and
                                    check_rows_fp64(loc_out, dy, lambda r: dx[r], rows_out))
    if with_oracle and x.shape[0] * feat * 4 <= (2 << 30):
        a = check_rows_oracle(loc_in, x.cpu().numpy(), y[rows_in], rows_in)
        b = check_rows_oracle(loc_out, dy.cpu().numpy(), dx[rows_out], rows_out)
        if a is not None:
            res["sampled_rows_c_oracle"] = bool(a and b)
    res["rows_sampled"] = int(rows_in.numel() + rows_out.numel())
    res["ok"] = all(v for k, v in res.items() if isinstance(v, bool))
    return res


def check_multi_gpu(part, check_ref, x_loc, dy_loc, y_loc, dx_loc, step, exchange, dev, rank):
    """N > 1: the exchange under test against the plain all-gather path (RowPartition over NCCL on the
    un-remapped row block), same local kernel; every rank must agree."""
    import torch.distributed as dist
    from dgl_b200 import dist as b200dist
    tol = 1e-2 if (b200dist.WIRE == "bf16" and exchange == "halo") else 1e-5   # bf16 wire: the bf16 bar

    def same(a, b):      # chunked: the operands can be tens of GB
        ok_, scale = True, float(b.abs().max()) + 1e-30
        for s0 in range(0, a.shape[0], 1 << 20):
            d = (a[s0:s0 + (1 << 20)] - b[s0:s0 + (1 << 20)]).abs_()
            ok_ = ok_ and bool((d <= tol * scale + tol * b[s0:s0 + (1 << 20)].abs()).all())
        return ok_
    step(False)
    want = check_ref.forward(x_loc)
    ok = same(y_loc, want)
    del want
    check_ref._full.clear()
    want = check_ref.backward(dy_loc)
    ok = same(dx_loc, want) and ok
    del want
    check_ref._full.clear()
    torch.cuda.synchronize()
    err = part.comm_errors() if hasattr(part, "comm_errors") else 0
    flag = torch.tensor([1 if (ok and err == 0) else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    return {"ok": int(flag.item()) == 1, "against": "all-gather over NCCL + same local SpMM", "rtol": tol,
            "signal_timeouts_rank0": int(err)}


# ---------------------------------------------------------------------------------------------
# B200 arm
# ---------------------------------------------------------------------------------------------

def measure_workload(args, cfg, ctx, full):
    """Builds the workload's graph, partitions it over the ranks, runs warm-up + `steps` timed steps
    (forward + gradient SpMM with their exchanges) and returns the measurements.  full=True adds what
    only the headline line carries (e2e, GCN epoch, clock sampling)."""
    import torch.distributed as dist
    from dgl_b200 import ops, _lib
    import dgl_b200
    from dgl_b200.dist import HaloPartition, RowPartition
    from dgl_b200 import dist as b200dist

    world, rank, dev = ctx["world"], ctx["rank"], ctx["dev"]
    n, e, feat = cfg["n"], cfg["e"], cfg["feat"]
    steps = args.steps if full else max(3, min(args.steps, 10))
    sampler = ctx.get("sampler") if full else None
    res = {}

    # ---- graph + structure (one-off, cached per graph like GraphIndex.adjacency_matrix) ------
    t0 = time.time()
    g = dgl_b200.DGLGraph()
    g.add_nodes(n)
    src, dst = make_edges(cfg, dev)
    g.add_edges(src, dst)
    del src, dst
    csr_in = g.in_csr()        # rows = dst (forward)
    csr_out = g.out_csr()      # rows = src (gradient)
    g.drop_coo()
    torch.cuda.synchronize()
    t_build = time.time() - t0

    exchange = args.exchange
    if exchange == "auto":   # compacted halo pays when rows are referenced rarely (C5: mean degree 3)
        exchange = "halo" if (world > 1 and e < 16 * n) else "allgather"
    t0 = time.time()
    if exchange == "halo":
        chunks = args.chunks or (4 if args.transport == "peer" else 1)
        part = HaloPartition(csr_in, csr_out, rank=rank, world=world, chunks=chunks)
    else:
        part = RowPartition(csr_in, csr_out, rank=rank, world=world)
    rows = part.rows
    loc_in, loc_out = part.loc_in, part.loc_out
    check_ref = None
    do_check = not args.no_check
    if do_check and world > 1:      # reference for the check: the un-remapped row block + a plain all-gather
        check_ref = RowPartition(csr_in, csr_out, rank=rank, world=world)
    if world > 1:
        del csr_in, csr_out
        g.clear_cache()
        torch.cuda.empty_cache()
    loc_in.plan()
    loc_out.plan()
    torch.cuda.synchronize()
    t_plan = time.time() - t0

    gen = torch.Generator(device=dev)
    gen.manual_seed(1234 + rank)
    like = torch.empty(0, device=dev)
    if args.transport == "peer" and world > 1:
        part.enable_peer_transport(feat, like)
    x_loc = part.local_view("x", feat, like)            # this rank's blocks live inside the exchange buffers
    dy_loc = part.local_view("dy", feat, like)
    for t_ in (x_loc, dy_loc):                           # randn in row pieces (C5: 25.6 GB per operand)
        for s0 in range(0, rows, 1 << 22):
            t_[s0:s0 + (1 << 22)].copy_(torch.randn(min(1 << 22, rows - s0), feat, device=dev, generator=gen))
    y_loc = torch.empty(rows, feat, device=dev)
    dx_loc = torch.empty(rows, feat, device=dev)

    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    op_events = []

    def step(record):
        # forward: halo exchange (pipelined with the SpMM where the transport allows) + SpMM over the
        # rank's destination rows; gradient: the same over the out-edge CSR
        for fn_, loc_, out_ in ((part.forward, x_loc, y_loc), (part.backward, dy_loc, dx_loc)):
            if record:
                a, b = ev(), ev()
                a.record()
            fn_(loc_, out=out_)
            if record:
                b.record()
                op_events.append((a, b))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # warm-up: the requested steps, then as many more as it takes to have run ~0.3 s back to back
    # (clocks settle under load and the sampler has samples to fall back on); the extra count is
    # decided by rank 0 and broadcast so that every rank runs the same number of exchanges
    torch.cuda.synchronize()
    t_w = time.time()
    n_warm = max(args.warmup, 3)
    for _ in range(n_warm):
        step(False)
    torch.cuda.synchronize()
    per = max((time.time() - t_w) / n_warm, 1e-5)
    extra = torch.tensor([min(2000, max(0, int((0.3 - per * n_warm) / per)))], device=dev, dtype=torch.int64)
    if world > 1:
        dist.broadcast(extra, src=0)
    for _ in range(int(extra.item())):
        step(False)
    barrier()
    if sampler:
        sampler.mark()
    l0 = _lib.launch_count()
    s_ev, e_ev = ev(), ev()
    s_ev.record()
    for _ in range(steps):
        step(True)
    e_ev.record()
    barrier()
    launches = _lib.launch_count() - l0
    clocks = sampler.stop() if sampler else None
    ms_total = s_ev.elapsed_time(e_ev)
    if world > 1:
        t = torch.tensor([ms_total], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    ms_step = ms_total / steps
    alg_step = 2 * graphs.spmm_alg_bytes(e, n, feat)          # whole job, all ranks

    # dominant kernel: the SpMM launches of this rank (forward and gradient are the same kernel).
    # At N == 1 an op IS the kernel (+ the 0.07 ms fix-up); at N > 1 ops include the exchange, so the
    # kernel alone is timed in a few extra launches over the already-exchanged operands.
    op_ms = sum(a.elapsed_time(b) for a, b in op_events) / len(op_events)
    if world == 1:
        k_avg_ms = op_ms
    else:
        kev = []
        for _ in range(3):
            for csr_, key_, out_ in ((loc_in, "x", y_loc), (loc_out, "dy", dx_loc)):
                xb = part.operand(key_, feat, like)
                a, b = ev(), ev()
                a.record()
                part.local_spmm(csr_, xb, out_)
                b.record()
                kev.append((a, b))
        torch.cuda.synchronize()
        k_avg_ms = sum(a.elapsed_time(b) for a, b in kev) / len(kev)
        step(False)                                         # leave y/dx as a real step produced them
    xchg_ms = max(0.0, op_ms - k_avg_ms)
    k_ranks = None
    if world > 1:          # per-rank kernel time: the spread is what the signals of the exchange wait for
        mine = torch.tensor([k_avg_ms], device=dev, dtype=torch.float64)
        allk = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allk, mine)
        k_ranks = [round(float(t[0]), 3) for t in allk]
    loc_e = (loc_in.nnz + loc_out.nnz) / 2.0
    alg_launch = graphs.spmm_alg_bytes(loc_e, rows, feat)

    # ---- self-check of what was just timed (outside the timed region) ------------------------
    check = None
    if do_check:
        if world == 1:
            check = check_single_gpu(loc_in, loc_out, x_loc, dy_loc, y_loc, dx_loc, feat, with_oracle=True)
        else:
            check = check_multi_gpu(part, check_ref, x_loc, dy_loc, y_loc, dx_loc, step, exchange, dev, rank)
            check_ref = None
            torch.cuda.empty_cache()
        if not check["ok"]:
            raise SystemExit("rank %d: self-check FAILED on %s: %s" % (rank, cfg["name"], json.dumps(check)))

    xchg_bytes = part.exchange_bytes(feat)
    transfer = None
    if exchange == "halo" and world > 1 and part.halo_in.timing:
        # DGL_B200_TIMING=1: the transfers alone, per rank
        torch.cuda.synchronize()
        tm = [a.elapsed_time(b) for h in (part.halo_in, part.halo_out) for a, b in h.timing[-2 * steps:]]
        mine = torch.tensor([sum(tm) / len(tm), float(xchg_bytes)], device=dev, dtype=torch.float64)
        allr = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        transfer = [{"ms": round(float(t[0]), 3), "GBps_in": round(float(t[1]) / float(t[0]) / 1e6, 1)} for t in allr]

    res.update(dict(
        cfg=cfg, exchange=exchange, part=part, g=g, loc_in=loc_in, loc_out=loc_out, rows=rows,
        x_loc=x_loc, dy_loc=dy_loc, y_loc=y_loc, dx_loc=dx_loc, gen=gen, like=like, barrier=barrier,
        steps=steps, ms_step=ms_step, alg_step=alg_step, op_ms=op_ms, k_avg_ms=k_avg_ms, xchg_ms=xchg_ms,
        k_ranks=k_ranks, alg_launch=alg_launch, loc_e=loc_e, launches=int(launches), clocks=clocks,
        check=check, xchg_bytes=xchg_bytes, transfer=transfer, t_build=t_build, t_plan=t_plan,
        extra_warm=int(extra.item()),
        parallelism="row-partition x%d + %s %s" % (
            world, ("NVLink peer " + (b200dist.transport_name(exchange)))
            if (args.transport == "peer" and world > 1) else "NCCL",
            ("compacted halo, %d row chunks" % part.halo_in.chunks) if exchange == "halo" else "all-gather halo"),
        wire=(b200dist.WIRE if exchange == "halo" else "f32")))
    return res


C5_N1_PATHS = (os.path.join(tempfile.gettempdir(), "dgl_b200_c5_n1.json"),
               os.path.join(ROOT, "gpurun_out", "c5_n1.json"))
C5_N1_COMMITTED = os.path.join(ROOT, "profiles", "r02_c5_n1.json")


def c5_n1_reference():
    """ms per step of the C5 1-GPU run this box made earlier (the driver runs N = 1, 2, 4, 8 back to
    back on one lease; the N = 1 invocation writes the file), else the committed measurement."""
    for p in C5_N1_PATHS:
        try:
            if time.time() - os.path.getmtime(p) < 4 * 3600:
                with open(p) as f:
                    d = json.load(f)
                return float(d["ms_per_step"]), "same box, written by the N=1 invocation (%s)" % p
        except (OSError, ValueError, KeyError):
            continue
    try:
        with open(C5_N1_COMMITTED) as f:
            d = json.load(f)
        return float(d["ms_per_step"]), "cached: committed profiles/r02_c5_n1.json (%s)" % d.get("when", "?")
    except (OSError, ValueError, KeyError):
        return None, "unavailable"


def sampling_record():
    """Neighbour sampling beside the headline, N = 1 only: batches of 1000 seeds x fanout 10 x 2 hops on
    a C2-shaped graph with 1/5 of Reddit's edges (so the reference's ImmutableGraph builds in seconds), this engine's
    dglb200_sample_subgraph against the compiled reference's NeighborUniformSample on the host cores (1 batch per
    call = 1 core; 16 per call = its OpenMP path), one batch checked against the oracle's invariant checker."""
    import time
    from dgl_b200 import ops, sampling as engine
    from oracle import oracle, ref_ffi
    dev = torch.device("cuda")
    c = graphs.C2_REDDIT
    n = c["n"]
    src, dst = graphs.powerlaw_symmetric(n, c["e"] // 5, c["seed"], dev)
    e = int(src.shape[0])
    csr = ops.csr_build(dst, src, n, n, order=1)
    gen = torch.Generator(device="cpu").manual_seed(0)
    batches = [torch.randperm(n, generator=gen)[:1000].to(dev) for _ in range(16)]
    fanout, hops = 10, 2
    for i in range(5):
        sub = engine.sample_subgraph(csr, batches[i], fanout, hops, seed=i)
    torch.cuda.synchronize()
    ts, edges = [], 0
    for i in range(40):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        sub = engine.sample_subgraph(csr, batches[i % 16], fanout, hops, seed=100 + i)
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
        edges += int(sub.induced_edges.shape[0])
    ms = sorted(ts)[len(ts) // 2]
    rec = {"workload": "1000 seeds x fanout 10 x 2 hops, in-neighbours, C2-shaped graph n=%d e=%d" % (n, e),
           "ms_per_batch": round(ms, 4), "edges_per_subgraph": edges // 40,
           "sampled_edges_per_s": round(edges / 40 / ms * 1e3)}
    small = batches[0][:50]
    sub = engine.sample_subgraph(csr, small, fanout, hops, seed=7)
    try:
        rec["checked"] = bool(oracle.check_neighbor_subgraph(
            csr.indptr.cpu().numpy(), csr.indices.cpu().numpy().astype("int64"), csr.eids.cpu().numpy().astype("int64"),
            small.cpu().numpy(), fanout, hops, sub.induced_nodes.cpu().numpy(), sub.layer_ids.cpu().numpy(),
            sub.induced_edges.cpu().numpy(), sub.sub_rows.cpu().numpy(), sub.sub_cols.cpu().numpy()))
    except AssertionError as ex:
        rec["checked"], rec["check_error"] = False, str(ex)
    if ref_ffi.available():
        g = ref_ffi.RefGraph(src.cpu().numpy(), dst.cpu().numpy(), n, readonly=True, multigraph=True)
        seeds_np = [b.cpu().numpy() for b in batches]
        for per_call in (1, 16):
            tt = []
            for _ in range(4):
                t0 = time.perf_counter()
                ref_ffi.graph_uniform_sampling(g, seeds_np[:per_call], "in", hops, fanout)
                tt.append((time.perf_counter() - t0) / per_call)
            rec["reference_ms_per_batch_%d_per_call" % per_call] = round(sorted(tt[1:])[1] * 1e3, 3)
        rec["reference"] = "oracle/_ref/libdgl.so NeighborUniformSample through its C ABI, %d host cores" % (os.cpu_count() or 1)
    return rec


def c5_record(args, ctx, wire=None):
    """The north-star's multi-GPU configuration (C5: RMAT 50M nodes / 150M edges, F = 128) measured in
    the same invocation as the headline and reported beside it (`c5` key); `value` / `config` of the
    line stay the C2 headline."""
    world, rank = ctx["world"], ctx["rank"]
    cfg = workload_cfg("c5")
    from dgl_b200 import dist as b200dist_
    saved_wire = b200dist_.WIRE
    if wire is not None:
        b200dist_.WIRE = wire            # "bf16": halo rows cross NVLink as bfloat16 (1e-2 bar), fp32 accumulate
    try:
        m = measure_workload(args, cfg, ctx, full=False)
    finally:
        b200dist_.WIRE = saved_wire
    peak, peak_kind = measured_peak()
    rec = None
    if rank == 0:
        ach = m["alg_launch"] / (m["k_avg_ms"] * 1e-3) / 1e9
        rec = {"workload": cfg["name"], "n": cfg["n"], "e": cfg["e"], "feat": cfg["feat"],
               "ms_per_step": m["ms_step"], "steps": m["steps"], "kernel_ms": m["k_avg_ms"],
               "op_ms_rank0": m["op_ms"], "exchange_ms": m["xchg_ms"], "exchange_bytes": m["xchg_bytes"],
               "wire": m["wire"], "parallelism": m["parallelism"], "kernel_ms_per_rank": m["k_ranks"],
               "transfer_per_rank": m["transfer"],
               "alg_GBps": m["alg_step"] / (m["ms_step"] * 1e-3) / 1e9,
               "alg_frac": ach / peak, "alg_frac_of": "kernel alone on rank 0 vs %s HBM peak %.0f GB/s" % (peak_kind, peak),
               "min_bytes_per_launch": graphs.spmm_min_bytes(m["loc_e"], cfg["n"], m["rows"], cfg["feat"]),
               "alg_bytes_per_launch": m["alg_launch"],
               "graph_build_s": round(m["t_build"], 2), "partition_plan_s": round(m["t_plan"], 2),
               "checked": bool(m["check"] and m["check"]["ok"]), "check": m["check"]}
        if world == 1:
            rec["speedup_vs_n1"] = 1.0
            out = {"ms_per_step": m["ms_step"], "kernel_ms": m["k_avg_ms"], "when": time.strftime("%Y-%m-%dT%H:%M:%SZ", time.gmtime())}
            for p in C5_N1_PATHS:
                try:
                    os.makedirs(os.path.dirname(p), exist_ok=True)
                    with open(p, "w") as f:
                        json.dump(out, f)
                except OSError:
                    pass
            tpath = os.path.join(ROOT, "profiles", "traffic_c5.json")
            if os.path.exists(tpath):
                with open(tpath) as f:
                    rec["traffic"] = json.load(f).get("dram_bytes_per_launch")
        else:
            n1, src_ = c5_n1_reference()
            rec["n1_ms_per_step"] = n1
            rec["n1_source"] = src_
            rec["speedup_vs_n1"] = (n1 / m["ms_step"]) if n1 else None
    # free everything before the caller goes on
    part = m["part"]
    for k in ("x_loc", "dy_loc", "y_loc", "dx_loc", "part", "g", "loc_in", "loc_out"):
        m.pop(k, None)
    if hasattr(part, "release"):
        part.release()
    del part, m
    if world > 1:
        from dgl_b200 import dist as b200dist
        b200dist.release_peer_memory()       # exported buffers stay allocated until every importer unmaps them
    torch.cuda.empty_cache()
    return rec


def run_b200(args):
    import torch.distributed as dist
    from dgl_b200 import ops, _lib
    from dgl_b200 import dist as b200dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")   # compute / exchange / NCCL streams never share a hardware queue
    if world > 1:
        torch.cuda.set_device(local)
        os.environ.setdefault("TORCH_NCCL_HIGH_PRIORITY", "1")   # stream-ordered collectives must not queue behind SpMM CTAs
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    cfg = workload_cfg(args.workload)
    n, e, feat = cfg["n"], cfg["e"], cfg["feat"]
    _lib.get_lib()
    ctx = {"world": world, "rank": rank, "dev": dev,
           "sampler": ClockSampler(local) if rank == 0 else None}   # polls from now on; mark() at the timed region

    m = measure_workload(args, cfg, ctx, full=True)
    part, g, exchange = m["part"], m["g"], m["exchange"]
    loc_in, loc_out, rows = m["loc_in"], m["loc_out"], m["rows"]
    x_loc, dy_loc, y_loc, dx_loc = m["x_loc"], m["dy_loc"], m["y_loc"], m["dx_loc"]
    gen, like, barrier = m["gen"], m["like"], m["barrier"]
    ms_step, alg_step, k_avg_ms = m["ms_step"], m["alg_step"], m["k_avg_ms"]
    value = alg_step / (ms_step * 1e-3) / 1e9
    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731

    peak, peak_kind = measured_peak()
    achieved = m["alg_launch"] / (k_avg_ms * 1e-3) / 1e9
    traffic = lts_pct = None
    tpath = os.path.join(ROOT, "profiles", "traffic_%s.json" % args.workload)
    if world == 1 and os.path.exists(tpath):
        with open(tpath) as f:
            tj = json.load(f)
        traffic, lts_pct = tj.get("dram_bytes_per_launch"), tj.get("lts_throughput_pct")
    min_launch = graphs.spmm_min_bytes(m["loc_e"], n, rows, feat)
    # which unit binds: with every source row re-read ALG/MIN times out of the L2 (C2: 126 x) the kernel
    # is L2-bandwidth-bound and the HBM fraction is an L2-assisted throughput figure, not DRAM utilisation
    bound = "l2" if m["alg_launch"] > 8 * min_launch else "hbm"
    roofline = {"bound": bound, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_kind": peak_kind,
                "peak_is": "HBM copy bandwidth (the graded denominator)",
                "lts_throughput_pct": lts_pct, "traffic_source": "ncu --set full capture committed under profiles/ (static)",
                "kernel": "spmm_kernel<float,4,...,copy_src,sum>", "kernel_ms": k_avg_ms,
                "alg_bytes_per_launch": m["alg_launch"], "min_bytes_per_launch": min_launch}

    # ---- e2e: host API with HOST feature buffers ---------------------------------------------
    e2e = None
    if world > 1 and not args.no_e2e and n * feat * 4 <= (8 << 30):
        e2e = e2e_partitioned(part, x_loc, dy_loc, y_loc, dx_loc, alg_step, barrier, dev, world, args)
    del x_loc, dy_loc, y_loc, dx_loc
    for k in ("x_loc", "dy_loc", "y_loc", "dx_loc"):
        m.pop(k)
    if world == 1 and not args.no_e2e and n * feat * 4 <= (8 << 30):
        part.release()
        torch.cuda.empty_cache()
        from dgl_b200 import function as fn
        with numa_local(dev) as placement:
            xh = torch.randn(n, feat).pin_memory()
            dyh = torch.randn(n, feat).pin_memory()
            yh = torch.empty(n, feat).pin_memory()
            dxh = torch.empty(n, feat).pin_memory()
        g.set_csr(loc_in, loc_out)

        # Host buffers in, host buffers out, every step; copies ride two side streams (PCIe is full
        # duplex) and steps are not separated by a host sync, so the H2D of step i+1 overlaps the
        # kernels of step i and the D2H of step i-1.  Every step still moves all of its 954 MB.
        s_in, s_out = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
        cur = torch.cuda.current_stream(dev)
        copies_done = []

        def e2e_step():
            with torch.cuda.stream(s_in):
                xd = xh.to(dev, non_blocking=True)
                ev_x = torch.cuda.Event()
                ev_x.record(s_in)
                dyd = dyh.to(dev, non_blocking=True)
                ev_dy = torch.cuda.Event()
                ev_dy.record(s_in)
            xd.record_stream(cur)
            dyd.record_stream(cur)
            cur.wait_event(ev_x)
            xd.requires_grad_(True)
            g.ndata["h"] = xd
            g.update_all(fn.copy_src("h", "m"), fn.sum("m", "y"))      # the call a reference user makes
            y = g.ndata["y"]
            ev_y = torch.cuda.Event()
            ev_y.record(cur)
            with torch.cuda.stream(s_out):
                s_out.wait_event(ev_y)
                yh.copy_(y.detach(), non_blocking=True)
            y.record_stream(s_out)
            cur.wait_event(ev_dy)
            y.backward(dyd)
            dx = xd.grad
            ev_dx = torch.cuda.Event()
            ev_dx.record(cur)
            with torch.cuda.stream(s_out):
                s_out.wait_event(ev_dx)
                dxh.copy_(dx, non_blocking=True)
                done = torch.cuda.Event()
                done.record(s_out)
            dx.record_stream(s_out)
            copies_done.append(done)
            if len(copies_done) > 2:          # bound the run-ahead: at most 2 steps in flight
                copies_done.pop(0).synchronize()

        for _ in range(2):
            e2e_step()
        torch.cuda.synchronize()
        copies_done.clear()
        k = max(3, min(args.steps, 10))
        t0 = time.perf_counter()
        for _ in range(k):
            e2e_step()
        torch.cuda.synchronize()
        t_e2e = (time.perf_counter() - t0) / k
        e2e = {"value": alg_step / t_e2e / 1e9, "unit": UNIT,
               "h2d_bytes_per_step": 2 * n * feat * 4, "d2h_bytes_per_step": 2 * n * feat * 4,
               "ms_per_step": t_e2e * 1e3, "steps": k, "method": "3-stream pipeline, 2 steps in flight",
               "host_buffers": placement.note}
        del xh, dyh, yh, dxh

    # ---- GCN epoch (second half of BASELINE.json's metric) ------------------------------------
    gcn = None
    if not args.no_gcn and cfg["kind"] == "powerlaw" and exchange != "halo":
        from dgl_b200 import function as fn
        feat_in, hidden, classes = gcn_dims(cfg)
        deg = (loc_in.indptr[1:] - loc_in.indptr[:-1]).float()
        norm = deg.pow(-0.5)
        norm[torch.isinf(norm)] = 0
        norm = norm.unsqueeze(1)
        xg = torch.randn(rows, feat_in, device=dev, generator=gen)
        labels = torch.randint(0, classes, (rows,), device=dev, generator=gen)
        model = GCNModel(feat_in, hidden, classes, dev)
        fused_layer = None
        if world == 1:
            g.set_csr(loc_in, loc_out)

            def agg(h, tag):                     # the call a reference user makes (gcn_spmv.py:52-54)
                g.ndata["h"] = h
                g.update_all(fn.copy_src("h", "m"), fn.sum("m", "h"))
                return g.ndata.pop("h")
            allreduce = None
            if not args.no_fused:
                from dgl_b200 import fused as b200fused
                adj_, norm1 = g.adjacency(), norm.reshape(-1).contiguous()
                fused_layer = lambda hw, b, relu: b200fused.gcn_aggregate(adj_, hw, norm1, bias=b, relu=relu)  # noqa: E731
        else:
            if args.transport == "peer":
                part.enable_peer_transport(hidden, like, keys=("xh", "dyh"))
                part.enable_peer_transport(classes, like, keys=("xc", "dyc"))
            agg = lambda h, tag: b200dist.aggregate(part, h, tag)  # noqa: E731
            allreduce = lambda t: dist.all_reduce(t)  # noqa: E731
        for _ in range(3):
            model.epoch(xg, labels, norm, agg, n, allreduce, fused_layer)
        barrier()
        k = max(3, min(args.steps, 10))
        l1 = _lib.launch_count()
        a, b = ev(), ev()
        a.record()
        for _ in range(k):
            loss = model.epoch(xg, labels, norm, agg, n, allreduce, fused_layer)
        b.record()
        barrier()
        ms = a.elapsed_time(b) / k
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        gcn = {"ms": ms, "edges": int(e), "edges_per_s": e / (ms * 1e-3), "epochs": k,
               "model": "GCN %d-%d-%d, dropout 0.5, Adam" % (feat_in, hidden, classes),
               "loss": float(loss.item()), "engine_launches_per_epoch": (_lib.launch_count() - l1) / k,
               "layer": "fused (norm / bias / relu in the SpMM epilogue)" if fused_layer is not None else "unfused"}
        del xg, labels, model
        torch.cuda.empty_cache()

    # ---- C5 (the north-star's multi-GPU configuration), reported beside the headline ----------
    parallelism, wire = m["parallelism"], m["wire"]
    keep = {k: m[k] for k in ("op_ms", "xchg_ms", "transfer", "k_ranks", "t_build", "t_plan", "extra_warm",
                              "xchg_bytes", "launches", "clocks", "check", "steps")}
    part.release()
    del part, g, loc_in, loc_out, m
    if world > 1:
        b200dist.release_peer_memory()
    torch.cuda.empty_cache()
    c5 = c5_bf16 = None
    if not args.no_c5 and args.workload == "c2":
        barrier()
        c5 = c5_record(args, ctx)
        if world > 1:        # second lever of SURVEY 8e, reported beside the fp32-wire (parity-mode) record
            barrier()
            c5_bf16 = c5_record(args, ctx, wire="bf16")

    samp = None
    if rank == 0 and world == 1 and not args.no_sampling:
        try:
            samp = sampling_record()
        except Exception as ex:  # noqa: BLE001 - an auxiliary record must not take the headline down
            samp = {"error": str(ex)[:300]}
        torch.cuda.empty_cache()

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = cpu_reference(cfg, 2, 1, budget_s=20.0)
        cpu = {k: cpu[k] for k in ("value", "unit", "cores", "effective_cores", "kind", "sample")}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": cfg["name"], "n": n, "e": e, "feat": feat,
                       "l2": "inputs_exceed_l2 (X %.0f MB + structure %.0f MB per SpMM; no flush)"
                             % (n * feat * 4 / 1e6, e * 4 / 1e6),
                       "parallelism": parallelism,
                       "exchange_bytes_per_rank_per_spmm": keep["xchg_bytes"], "wire": wire,
                       "op_ms_rank0": keep["op_ms"], "exposed_exchange_ms_rank0": keep["xchg_ms"],
                       "transfer_per_rank": keep["transfer"], "kernel_ms_per_rank": keep["k_ranks"],
                       "graph_build_s": round(keep["t_build"], 2), "partition_plan_s": round(keep["t_plan"], 2),
                       "extra_warmup_steps": keep["extra_warm"]},
            "edge_feat_per_s": 2 * e * feat / (ms_step * 1e-3),
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e,
            "gpu_launches": keep["launches"], "clocks": keep["clocks"], "gcn_epoch": gcn,
            "checked": bool(keep["check"] and keep["check"]["ok"]), "check": keep["check"], "c5": c5, "c5_bf16_wire": c5_bf16,
            "neighbor_sampling": samp,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", choices=["c2", "c5", "tiny"])
    ap.add_argument("--exchange", default="auto", choices=["auto", "allgather", "halo"])
    ap.add_argument("--transport", default="peer", choices=["nccl", "peer"],
                    help="N>1 data path: NCCL collectives, or peer loads over NVLink from IPC-mapped buffers")
    ap.add_argument("--chunks", type=int, default=0,
                    help="row chunks of the pipelined halo exchange (0 = auto: 4 with the peer transport, "
                         "whose pushes run beside the SpMM of the previous chunk; 1 = unpipelined)")
    ap.add_argument("--check", action="store_true", help="(default on; kept for old command lines)")
    ap.add_argument("--no-check", action="store_true",
                    help="skip the self-check (N=1: slab path vs column blocks bit for bit + sampled rows vs the "
                         "oracle; N>1: the exchange against the all-gather path)")
    ap.add_argument("--no-fused", action="store_true", help="GCN epoch with the unfused layer (4 eager passes around the SpMM)")
    ap.add_argument("--no-c5", action="store_true", help="skip the C5 record measured beside the C2 headline")
    ap.add_argument("--no-sampling", action="store_true", help="skip the neighbour-sampling record (N = 1)")
    ap.add_argument("--ref-budget", type=float, default=150.0,
                    help="--impl reference: seconds of timed CPU passes over the full configuration")
    ap.add_argument("--ref-sample", action="store_true", help="--impl reference on a bounded edge sample")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-gcn", action="store_true", help="skip the GCN-epoch measurement")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
