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

def cpu_reference(cfg, steps, warmup, budget_s=25.0, gcn=False):
    """Times fwd+bwd of th.sparse.mm on a bounded sample (same N and F, E/k edges of the same
    generator) sized so that steps+warmup passes take about `budget_s` seconds."""
    from oracle import oracle
    n, feat = cfg["n"], cfg["feat"]
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
    total = max(1, steps + warmup)
    e_s = int(min(cfg["e"], max(probe_e, budget_s / total / per_edge)))
    e_s -= e_s % 2
    src, dst = make_edges(cfg, "cpu", e_s, e_s / cfg["e"])
    ts = one_pass(src, dst, total)[warmup:] or [t_probe]
    t = sum(ts) / len(ts)
    alg = 2 * graphs.spmm_alg_bytes(e_s, n, feat)
    gcn = gcn_epoch_cpu(src, dst, n, cfg) if (gcn and cfg["kind"] == "powerlaw") else None
    return dict(gcn_epoch=gcn, value=alg / t / 1e9, unit=UNIT, cores=cores, kind="port",
                sample="th.sparse.mm(COO)+autograd bwd, fp32, N=%d F=%d, E=%d of %d edges (1/%.1f), "
                       "%d timed passes, %.3f s/pass; %d torch threads offered, %.1f busy on average"
                       % (n, feat, e_s, cfg["e"], cfg["e"] / e_s,
                                                                  len(ts), t, cores, busy[0]),
                effective_cores=round(busy[0], 2),
                edge_feat_per_s=2 * e_s * feat / t, ms_per_step=t * 1e3, sample_edges=e_s)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg = workload_cfg(args.workload)
    r = cpu_reference(cfg, args.steps, args.warmup, budget_s=60.0, gcn=not args.no_gcn)
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": cfg["name"], "n": cfg["n"], "e": cfg["e"], "feat": cfg["feat"],
                   "sample_edges": r["sample_edges"]},
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

    def forward(self, x, norm, aggregate):
        drop = torch.nn.functional.dropout
        h = torch.mm(drop(x, self.dropout), self.w1) * norm
        h = torch.relu(aggregate(h, "h") * norm + self.b1)
        h = torch.mm(drop(h, self.dropout), self.w2) * norm
        return aggregate(h, "c") * norm + self.b2

    def epoch(self, x, labels, norm, aggregate, n_total, allreduce=None):
        logits = self.forward(x, norm, aggregate)
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


def b200dist_mode():
    from dgl_b200 import dist as b200dist
    return b200dist.PULL_MODE


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

def e2e_partitioned(part, x_loc, dy_loc, y_loc, dx_loc, alg_step, barrier, dev, world, args):
    """N > 1 end-to-end: every rank's row blocks of X and dY start in HOST (pinned) memory and its
    blocks of Y and dX end there, every step, through the partition's public calls
    (RowPartition / HaloPartition .forward / .backward = exchange + SpMM).  All copies and kernels
    are stream-ordered on the current stream; the exchange's own barriers keep a block from being
    overwritten while a peer still reads it.  Timed with CUDA events, max over ranks."""
    import torch.distributed as dist
    rows, feat = x_loc.shape
    xh = torch.randn(rows, feat).pin_memory()
    dyh = torch.randn(rows, feat).pin_memory()
    yh = torch.empty(rows, feat).pin_memory()
    dxh = torch.empty(rows, feat).pin_memory()

    def one():
        x_loc.copy_(xh, non_blocking=True)
        dy_loc.copy_(dyh, non_blocking=True)
        part.forward(x_loc, out=y_loc)
        yh.copy_(y_loc, non_blocking=True)
        part.backward(dy_loc, out=dx_loc)
        dxh.copy_(dx_loc, non_blocking=True)

    for _ in range(2):
        one()
    barrier()
    k = max(3, min(args.steps, 10))
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(k):
        one()
    b.record()
    barrier()
    t = torch.tensor([a.elapsed_time(b) / k], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    nbytes = torch.tensor([2.0 * rows * feat * 4], device=dev, dtype=torch.float64)
    dist.all_reduce(nbytes)                                   # whole job, all ranks
    return {"value": alg_step / (ms * 1e-3) / 1e9, "unit": UNIT,
            "h2d_bytes_per_step": int(nbytes.item()), "d2h_bytes_per_step": int(nbytes.item()),
            "ms_per_step": ms, "steps": k}


def run_b200(args):
    import torch.distributed as dist
    from dgl_b200 import ops, _lib
    import dgl_b200

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        torch.cuda.set_device(local)
        os.environ.setdefault("TORCH_NCCL_HIGH_PRIORITY", "1")   # the stream-ordered barriers must not queue behind SpMM CTAs
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    cfg = workload_cfg(args.workload)
    n, e, feat = cfg["n"], cfg["e"], cfg["feat"]
    lib = _lib.get_lib()
    sampler = ClockSampler(local) if rank == 0 else None     # polls from now on; mark() at the timed region

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

    from dgl_b200.dist import HaloPartition, RowPartition
    exchange = args.exchange
    if exchange == "auto":   # compacted halo pays when rows are referenced rarely (C5: mean degree 3)
        exchange = "halo" if (world > 1 and e < 16 * n) else "allgather"
    if exchange == "halo":
        chunks = args.chunks or (4 if args.transport == "peer" else 1)
        part = HaloPartition(csr_in, csr_out, rank=rank, world=world, chunks=chunks)
    else:
        part = RowPartition(csr_in, csr_out, rank=rank, world=world)
    bounds, r0, r1, rows = part.bounds, part.r0, part.r1, part.rows
    loc_in, loc_out = part.loc_in, part.loc_out
    check_ref = None
    if args.check and world > 1:      # reference for --check: the un-remapped row block + a plain all-gather
        check_ref = RowPartition(csr_in, csr_out, rank=rank, world=world)
    if world > 1:
        del csr_in, csr_out
        g.clear_cache()
        torch.cuda.empty_cache()
    loc_in.plan()
    loc_out.plan()

    gen = torch.Generator(device=dev)
    gen.manual_seed(1234 + rank)
    like = torch.empty(0, device=dev)
    if args.transport == "peer" and world > 1:
        part.enable_peer_transport(feat, like)
    x_loc = part.local_view("x", feat, like)            # this rank's blocks live inside the exchange buffers
    dy_loc = part.local_view("dy", feat, like)
    x_loc.copy_(torch.randn(rows, feat, device=dev, generator=gen))
    dy_loc.copy_(torch.randn(rows, feat, device=dev, generator=gen))
    y_loc = torch.empty(rows, feat, device=dev)
    dx_loc = torch.empty(rows, feat, device=dev)

    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    kern_events = []

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
    for _ in range(args.steps):
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
    ms_step = ms_total / args.steps
    alg_step = 2 * graphs.spmm_alg_bytes(e, n, feat)          # whole job, all ranks
    value = alg_step / (ms_step * 1e-3) / 1e9

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
                ops.spmm(csr_, ops.MSG_COPY_SRC, ops.RED_SUM, x=xb, out=out_)
                b.record()
                kev.append((a, b))
        torch.cuda.synchronize()
        k_avg_ms = sum(a.elapsed_time(b) for a, b in kev) / len(kev)
    xchg_ms = max(0.0, op_ms - k_avg_ms)
    k_ranks = None
    if world > 1:          # per-rank kernel time: the spread is what the barriers of the exchange wait for
        mine = torch.tensor([k_avg_ms], device=dev, dtype=torch.float64)
        allk = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allk, mine)
        k_ranks = [round(float(t[0]), 3) for t in allk]
    loc_e = (loc_in.nnz + loc_out.nnz) / 2.0
    alg_launch = graphs.spmm_alg_bytes(loc_e, rows, feat)
    peak, peak_kind = measured_peak()
    achieved = alg_launch / (k_avg_ms * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic_%s.json" % args.workload)
    if world == 1 and os.path.exists(tpath):
        with open(tpath) as f:
            traffic = json.load(f).get("dram_bytes_per_launch")
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_kind": peak_kind,
                "kernel": "spmm_kernel<float,4,...,copy_src,sum>", "kernel_ms": k_avg_ms,
                "alg_bytes_per_launch": alg_launch,
                "min_bytes_per_launch": graphs.spmm_min_bytes(loc_e, n, rows, feat)}

    # ---- e2e: host API with HOST feature buffers ---------------------------------------------
    e2e = None
    if check_ref is not None:
        # multi-GPU parity: the exchange under test vs the full all-gather path, same local kernel
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
        flag = torch.tensor([1 if ok else 0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) != 1:
            raise SystemExit("rank %d: --check FAILED: exchange '%s' disagrees with the all-gather path" % (rank, exchange))
        check_ref = None
        torch.cuda.empty_cache()
    xchg_bytes = part.exchange_bytes(feat)
    e2e_multi = None
    if world > 1 and not args.no_e2e and n * feat * 4 <= (8 << 30):
        e2e_multi = e2e_partitioned(part, x_loc, dy_loc, y_loc, dx_loc, alg_step, barrier, dev, world, args)
    transfer = None
    if exchange == "halo" and world > 1 and part.halo_in.timing:
        # DGL_B200_TIMING=1: the transfers alone (between the barriers), per rank
        torch.cuda.synchronize()
        tm = [a.elapsed_time(b) for h in (part.halo_in, part.halo_out) for a, b in h.timing[-2 * args.steps:]]
        mine = torch.tensor([sum(tm) / len(tm), float(xchg_bytes)], device=dev, dtype=torch.float64)
        allr = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        transfer = [{"ms": round(float(t[0]), 3), "GBps_in": round(float(t[1]) / float(t[0]) / 1e6, 1)} for t in allr]
    del x_loc, dy_loc, y_loc, dx_loc
    part._full.clear()
    if exchange == "halo":
        part.halo_in._buf = part.halo_out._buf = None
    torch.cuda.empty_cache()
    if world == 1 and not args.no_e2e and n * feat * 4 <= (8 << 30):
        from dgl_b200 import function as fn
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
               "ms_per_step": t_e2e * 1e3, "steps": k}
        del xh, dyh, yh, dxh
    elif world > 1:
        e2e = e2e_multi

    # ---- GCN epoch (second half of BASELINE.json's metric) ------------------------------------
    gcn = None
    if not args.no_gcn and cfg["kind"] == "powerlaw" and exchange != "halo":
        from dgl_b200 import function as fn
        from dgl_b200 import dist as b200dist
        feat_in, hidden, classes = gcn_dims(cfg)
        deg = (loc_in.indptr[1:] - loc_in.indptr[:-1]).float()
        norm = deg.pow(-0.5)
        norm[torch.isinf(norm)] = 0
        norm = norm.unsqueeze(1)
        xg = torch.randn(rows, feat_in, device=dev, generator=gen)
        labels = torch.randint(0, classes, (rows,), device=dev, generator=gen)
        model = GCNModel(feat_in, hidden, classes, dev)
        if world == 1:
            g.set_csr(loc_in, loc_out)

            def agg(h, tag):                     # the call a reference user makes (gcn_spmv.py:52-54)
                g.ndata["h"] = h
                g.update_all(fn.copy_src("h", "m"), fn.sum("m", "h"))
                return g.ndata.pop("h")
            allreduce = None
        else:
            if args.transport == "peer":
                part.enable_peer_transport(hidden, like, keys=("xh", "dyh"))
                part.enable_peer_transport(classes, like, keys=("xc", "dyc"))
            agg = lambda h, tag: b200dist.aggregate(part, h, tag)  # noqa: E731
            allreduce = lambda t: dist.all_reduce(t)  # noqa: E731
        for _ in range(3):
            model.epoch(xg, labels, norm, agg, n, allreduce)
        barrier()
        k = max(3, min(args.steps, 10))
        l1 = _lib.launch_count()
        a, b = ev(), ev()
        a.record()
        for _ in range(k):
            loss = model.epoch(xg, labels, norm, agg, n, allreduce)
        b.record()
        barrier()
        ms = a.elapsed_time(b) / k
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        gcn = {"ms": ms, "edges": int(e), "edges_per_s": e / (ms * 1e-3), "epochs": k,
               "model": "GCN %d-%d-%d, dropout 0.5, Adam" % (feat_in, hidden, classes),
               "loss": float(loss.item()), "engine_launches_per_epoch": (_lib.launch_count() - l1) / k}
        del xg, labels, model
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
                       "parallelism": "row-partition x%d + %s %s" % (
                           world, ("NVLink peer " + (b200dist_mode() if exchange == "halo" else "block copies"))
                           if (args.transport == "peer" and world > 1) else "NCCL",
                           ("compacted halo, %d row chunks" % part.halo_in.chunks) if exchange == "halo"
                           else "all-gather halo"),
                       "exchange_bytes_per_rank_per_spmm": xchg_bytes,
                       "wire": (os.environ.get("DGL_B200_WIRE", "f32") if exchange == "halo" else "f32"),
                       "op_ms_rank0": op_ms, "exposed_exchange_ms_rank0": xchg_ms, "transfer_per_rank": transfer, "kernel_ms_per_rank": k_ranks,
                       "graph_build_s": round(t_build, 2), "extra_warmup_steps": int(extra.item())},
            "edge_feat_per_s": 2 * e * feat / (ms_step * 1e-3),
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e,
            "gpu_launches": int(launches), "clocks": clocks, "gcn_epoch": gcn,
            "checked": bool(args.check and world > 1),
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
    ap.add_argument("--check", action="store_true", help="N>1: verify the exchange against the all-gather path")
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
