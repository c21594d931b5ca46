"""Sweep of the C5 multi-GPU exchange knobs in ONE launch (the graph is built once per rank):
transport (signal / round-1 push), row chunks, wire precision, push grid.  torchrun-launched:

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
      --master-port 29511 scripts/c5_sweep.py [--workload c5] [--quick]

Rank 0 prints one JSON line per configuration (ms per step = forward + gradient, max over ranks;
kernel-only ms; per-rank push ms with DGL_B200_TIMING) and writes them to gpurun_out/c5_sweep_N.jsonl.
Every configuration is checked against the all-gather path once (fp32 wire: rtol 1e-5; bf16: 1e-2).
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "dgl-1_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
os.environ.setdefault("TORCH_NCCL_HIGH_PRIORITY", "1")

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import bench  # noqa: E402
from dgl_b200 import dist as dmod, ops  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c5")
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--steps", type=int, default=8)
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    cfg = bench.workload_cfg(args.workload)
    n, e, feat = cfg["n"], cfg["e"], cfg["feat"]
    src, dst = bench.make_edges(cfg, dev)
    csr_in = ops.csr_build(dst, src, n, n)
    csr_out = ops.csr_build(src, dst, n, n)
    del src, dst
    torch.cuda.empty_cache()
    like = torch.empty(0, device=dev)
    gen = torch.Generator(device=dev)
    gen.manual_seed(99 + rank)
    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    out_lines = []

    # reference result of the check: plain all-gather over NCCL + the local kernel
    ref = dmod.RowPartition(csr_in, csr_out, rank=rank, world=world)
    rows = ref.rows
    x_src = torch.randn(rows, feat, device=dev, generator=gen)
    dy_src = torch.randn(rows, feat, device=dev, generator=gen)
    want_y = ref.forward(x_src).clone()
    want_dx = ref.backward(dy_src).clone()
    ref.release()
    del ref
    torch.cuda.empty_cache()

    # (chunks, chunk fractions or None = default, transport, wire, push grid)
    if args.quick:
        grid = [(4, None, "signal", "f32", 0), (4, None, "signal", "bf16", 0)]
    else:
        grid = [(1, None, "signal", "f32", 0),
                (4, None, "signal", "f32", 0), (4, "0.25,0.5,0.75,1", "signal", "f32", 0),
                (4, "0.1,0.35,0.7,1", "signal", "f32", 0), (4, None, "signal", "f32", 592),
                (6, None, "signal", "f32", 0), (8, None, "signal", "f32", 0),
                (8, "0.05,0.15,0.3,0.45,0.6,0.75,0.9,1", "signal", "f32", 0),
                (1, None, "signal", "bf16", 0), (2, None, "signal", "bf16", 0),
                (4, None, "signal", "bf16", 0), (4, "0.1,0.35,0.7,1", "signal", "bf16", 0),
                (8, None, "signal", "bf16", 0),
                (4, None, "push", "f32", 0)]
    if os.environ.get("C5_SWEEP_GRID"):       # e.g. '[[4,null,"signal","f32",64],[4,null,"signal-lsu","bf16",0]]'
        grid = [tuple(g) for g in json.loads(os.environ["C5_SWEEP_GRID"])]
    part, part_key = None, None
    dmod.TIMING = True
    for chunks, fracs, mode, wire, ctas in grid:
        if part is not None:
            part.release()
            dmod.release_peer_memory()
        if part is None or part_key != (chunks, fracs):
            part = None
            torch.cuda.empty_cache()
            if fracs:
                os.environ["DGL_B200_CHUNK_FRACS"] = fracs
            else:
                os.environ.pop("DGL_B200_CHUNK_FRACS", None)
            t0 = time.time()
            part = dmod.HaloPartition(csr_in, csr_out, rank=rank, world=world, chunks=chunks)
            part.loc_in.plan()
            part.loc_out.plan()
            torch.cuda.synchronize()
            plan_s = time.time() - t0
            part_key = (chunks, fracs)
        torch.cuda.empty_cache()
        # "signal-lsu": every thread stores to the peer; "signal-tma1": thread gather + TMA store; "signal": TMA both ways
        dmod.PUSH_TMA = 0 if mode.endswith("-lsu") else (1 if mode.endswith("-tma1") else 2)
        mode = mode.replace("-lsu", "").replace("-tma1", "")
        dmod.apply_tuning()
        dmod.PULL_MODE, dmod.WIRE, dmod.PUSH_CTAS = mode, wire, ctas
        part.enable_peer_transport(feat, like)
        x_loc, dy_loc = part.local_view("x", feat, like), part.local_view("dy", feat, like)
        x_loc.copy_(x_src)
        dy_loc.copy_(dy_src)
        y = torch.empty(rows, feat, device=dev)
        dx = torch.empty(rows, feat, device=dev)
        for _ in range(3):
            part.forward(x_loc, out=y)
            part.backward(dy_loc, out=dx)
        tol = 1e-2 if wire == "bf16" else 1e-5
        ok = True
        for got, want in ((y, want_y), (dx, want_dx)):
            scale = float(want.abs().max()) + 1e-30
            for s0 in range(0, rows, 1 << 20):          # chunked: the operands can be tens of GB
                g_, w_ = got[s0:s0 + (1 << 20)], want[s0:s0 + (1 << 20)]
                ok = ok and bool(((g_ - w_).abs() <= tol * scale + tol * w_.abs()).all())
        ok = ok and part.comm_errors() == 0
        flag = torch.tensor([1 if ok else 0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        for h in (part.halo_in, part.halo_out):
            h.timing.clear()
            h.timeline.clear()
        dist.barrier()
        torch.cuda.synchronize()
        a, b = ev(), ev()
        a.record()
        for _ in range(args.steps):
            part.forward(x_loc, out=y)
            part.backward(dy_loc, out=dx)
        b.record()
        dist.barrier()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / args.steps
        # kernel alone over the exchanged operands
        ka, kb = ev(), ev()
        ka.record()
        for _ in range(3):
            part.local_spmm(part.loc_in, part.operand("x", feat, like), y)
            part.local_spmm(part.loc_out, part.operand("dy", feat, like), dx)
        kb.record()
        torch.cuda.synchronize()
        k_ms = ka.elapsed_time(kb) / 3
        # timeline of the last forward op on this rank (ms after the op's start on the compute stream):
        # push k [start, end] on the exchange stream; chunk k [halo landed, SpMM done] on the compute stream
        tl = None
        if part.halo_in.timeline:
            evs = part.halo_in.timeline[-1]
            pushes = part.halo_in.timing[-chunks:]
            tl = {"push": [[round(evs[0].elapsed_time(p), 3), round(evs[0].elapsed_time(q), 3)] for p, q in pushes],
                  "chunk": [[round(evs[0].elapsed_time(evs[1 + 2 * k]), 3), round(evs[0].elapsed_time(evs[2 + 2 * k]), 3)]
                            for k in range(chunks)]}
        tm = [p.elapsed_time(q) for h in (part.halo_in, part.halo_out) for p, q in h.timing]
        push_ms = sum(tm) / max(len(tm), 1) * chunks          # per SpMM op: all chunks
        t = torch.tensor([ms, k_ms, push_ms], device=dev, dtype=torch.float64)
        allt = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(allt, t)
        if rank == 0:
            rec = {"world": world, "workload": args.workload, "chunks": chunks, "mode": mode + {0: "-lsu", 1: "-tma1", 2: ""}[dmod.PUSH_TMA], "wire": wire,
                   "push_ctas": ctas, "ms_per_step": round(max(float(v[0]) for v in allt), 3),
                   "kernel_ms_per_step_by_rank": [round(float(v[1]), 3) for v in allt],
                   "push_ms_per_op_by_rank": [round(float(v[2]), 3) for v in allt],
                   "halo_bytes_per_rank_per_op": part.exchange_bytes(feat), "plan_s": round(plan_s, 1),
                   "chunk_fracs": fracs, "rows_per_chunk": [part.halo_in.row_cuts[k + 1] - part.halo_in.row_cuts[k] for k in range(chunks)],
                   "halo_rows_per_chunk": [sum(c) for c in part.halo_in.recv_counts],
                   "timeline_fwd_rank0_ms": tl, "checked": int(flag.item()) == 1}
            print(json.dumps(rec), flush=True)
            out_lines.append(rec)
        del x_loc, dy_loc, y, dx
    if rank == 0:
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", "c5_sweep_%d.jsonl" % world), "w") as f:
            for r in out_lines:
                f.write(json.dumps(r) + "\n")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
