"""Multi-GPU probe (torchrun): where does the C5 halo exchange time go?  Times barrier, halo fetch
(full grid vs bounded grid), chunk SpMMs and the pipelined op separately on every rank."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "dgl-1_b200"))
import bench  # noqa: E402
import dgl_b200  # noqa: E402
from dgl_b200 import dist as bdist, ops  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
cfg = bench.workload_cfg(sys.argv[1] if len(sys.argv) > 1 else "c5")
chunks = int(sys.argv[2]) if len(sys.argv) > 2 else 4
n, feat = cfg["n"], cfg["feat"]
g = dgl_b200.DGLGraph()
g.add_nodes(n)
src, dst = bench.make_edges(cfg, dev)
g.add_edges(src, dst)
del src, dst
cin, cout = g.in_csr(), g.out_csr()
g.drop_coo()
part = bdist.HaloPartition(cin, cout, rank=rank, world=world, chunks=chunks)
del cin, cout
g.clear_cache()
like = torch.empty(0, device=dev)
part.enable_peer_transport(feat, like, keys=("x",))
h = part.halo_in
x_loc = part.local_view("x", feat, like)
x_loc.normal_()
y = torch.empty(part.rows, feat, device=dev)
buf = h.buffer(feat, like)
for c in h.row_chunks():
    c.plan()
h.csr.plan()


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    dist.barrier()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
        dist.barrier()
    return min(ts), sum(ts) / len(ts)


def fetch_all():
    h._barrier()
    for k in range(h.chunks):
        h._fetch_chunk(k, buf)
    h._barrier()


def spmm_chunks():
    for k, c in enumerate(h.row_chunks()):
        ops.spmm(c, ops.MSG_COPY_SRC, ops.RED_SUM, x=buf, out=y[h.row_cuts[k]:h.row_cuts[k + 1]])


res = {}
res["barrier_x2"] = timed(lambda: (h._barrier(), h._barrier()))
for ctas in (0, 256):
    bdist.PULL_CTAS = ctas
    res["fetch_ctas%d" % ctas] = timed(fetch_all)
res["spmm_whole"] = timed(lambda: ops.spmm(h.csr, ops.MSG_COPY_SRC, ops.RED_SUM, x=buf, out=y))
res["spmm_chunks"] = timed(spmm_chunks)
for prio in (-1,):
    bdist.COMM_PRIORITY = prio
    h._comm = None
    for ctas in (0, 128, 256, 512):
        bdist.PULL_CTAS = ctas
        res["pipelined_prio%d_ctas%d" % (prio, ctas)] = timed(lambda: h.aggregate(x_loc, part._spmm, y))
sizes = [sum(c) * feat * 4 / 1e9 for c in h.recv_counts]
for r in range(world):
    if r == rank:
        print("rank", rank, "rows", part.rows, "halo GB per chunk", [round(s, 3) for s in sizes],
              {k: (round(v[0], 3), round(v[1], 3)) for k, v in res.items()}, flush=True)
    dist.barrier()
dist.destroy_process_group()
