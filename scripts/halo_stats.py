"""GPU diagnostic for SURVEY.md 8e: how many DISTINCT remote source rows does each rank of a P-way
edge-balanced row partition need (the compacted-halo volume) vs the full all-gather volume."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "dgl-1_b200"))
import bench  # noqa: E402
import dgl_b200  # noqa: E402
from dgl_b200.dist import edge_balanced_bounds  # noqa: E402


def main():
    wl = sys.argv[1] if len(sys.argv) > 1 else "c5"
    cfg = bench.workload_cfg(wl)
    dev = torch.device("cuda")
    g = dgl_b200.DGLGraph()
    g.add_nodes(cfg["n"])
    src, dst = bench.make_edges(cfg, dev)
    g.add_edges(src, dst)
    del src, dst
    cin, cout = g.in_csr(), g.out_csr()
    g.drop_coo()
    n, feat = cfg["n"], cfg["feat"]
    deg_out = (cout.indptr[1:] - cout.indptr[:-1])
    out = {"workload": wl, "n": n, "nodes_with_out_edges": int((deg_out > 0).sum())}
    for parts in (2, 4, 8):
        b = edge_balanced_bounds(cin.indptr, cout.indptr, parts)
        rows = []
        for p in range(parts):
            lo, hi = int(cin.indptr[b[p]]), int(cin.indptr[b[p + 1]])
            cols = cin.indices[lo:hi]
            remote = cols[(cols < b[p]) | (cols >= b[p + 1])]
            uniq = torch.unique(remote)
            rows.append({"rank": p, "rows": b[p + 1] - b[p], "in_edges": hi - lo,
                         "remote_edges": int(remote.numel()), "unique_remote_src": int(uniq.numel()),
                         "full_gather_rows": n - (b[p + 1] - b[p])})
            del cols, remote, uniq
        tot_u = sum(r["unique_remote_src"] for r in rows)
        tot_f = sum(r["full_gather_rows"] for r in rows)
        out["P%d" % parts] = {"per_rank": rows, "compacted_over_full": tot_u / tot_f,
                              "compacted_GB_per_rank_fp32": max(r["unique_remote_src"] for r in rows) * feat * 4 / 1e9,
                              "full_GB_per_rank_fp32": max(r["full_gather_rows"] for r in rows) * feat * 4 / 1e9}
        print(parts, out["P%d" % parts]["compacted_over_full"], out["P%d" % parts]["compacted_GB_per_rank_fp32"],
              out["P%d" % parts]["full_GB_per_rank_fp32"], flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "halo_stats_%s.json" % wl), "w"), indent=1)


if __name__ == "__main__":
    main()
