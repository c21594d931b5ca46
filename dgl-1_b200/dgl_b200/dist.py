"""Multi-GPU message passing: 1-D destination-row partition + feature-block all-gather.

The reference has no distributed layer; this is the north-star's
scheme: one process per GPU, contiguous edge-balanced row blocks, rank p owns the rows
[bounds[p], bounds[p+1]) of every node tensor and of both CSRs.  Forward: all-gather the source
feature blocks (the halo exchange), then the local SpMM over the rank's destination rows.
Gradient: all-gather the dY blocks, then the local transposed SpMM over the rank's source rows
(gather form: no reduce-scatter, no atomics, deterministic).

Only torch.distributed is used for the exchange (NCCL on GPUs, gloo in the CPU tests); the
local compute is `ops.spmm` (libdgl_b200.so) unless a test injects a checker.
"""
import torch
import torch.distributed as dist

from . import ops


def edge_balanced_bounds(indptr_in, indptr_out, parts):
    """Row-block boundaries [0 = b_0 <= ... <= b_parts = n] such that every block holds about
    the same number of in + out edges (forward and gradient work)."""
    load = (indptr_in + indptr_out).to(torch.float64)
    n = int(indptr_in.shape[0] - 1)
    total = float(load[-1])
    if parts == 1 or total == 0:
        step = (n + parts - 1) // parts
        return [min(n, p * step) for p in range(parts)] + [n]
    targets = torch.arange(1, parts, dtype=torch.float64, device=load.device) * (total / parts)
    cuts = torch.searchsorted(load, targets).clamp_(max=n).tolist()
    bounds = [0] + [int(c) for c in cuts] + [n]
    for i in range(1, len(bounds)):            # monotone
        bounds[i] = max(bounds[i], bounds[i - 1])
    return bounds


class RowPartition(object):
    """This rank's share of a graph: row blocks of the in-edge CSR (forward) and of the out-edge
    CSR (gradient), plus the all-gather plumbing for node features."""

    def __init__(self, csr_in, csr_out, rank=None, world=None, group=None, spmm_fn=None):
        self.group = group
        self.world = world if world is not None else (dist.get_world_size(group) if dist.is_initialized() else 1)
        self.rank = rank if rank is not None else (dist.get_rank(group) if dist.is_initialized() else 0)
        self.n = csr_in.n_rows
        self.bounds = edge_balanced_bounds(csr_in.indptr, csr_out.indptr, self.world)
        self.r0, self.r1 = self.bounds[self.rank], self.bounds[self.rank + 1]
        self.rows = self.r1 - self.r0
        if self.world > 1:
            self.loc_in = ops.csr_row_block(csr_in, self.r0, self.r1)
            self.loc_out = ops.csr_row_block(csr_out, self.r0, self.r1)
        else:
            self.loc_in, self.loc_out = csr_in, csr_out
        self._spmm = spmm_fn or (lambda csr, x, out: ops.spmm(csr, ops.MSG_COPY_SRC, ops.RED_SUM, x=x, out=out)[0])
        self._full = {}

    def _global_rank(self, p):
        return p if self.group is None else dist.get_global_rank(self.group, p)

    def full_buffer(self, key, feat, like):
        """Persistent [n, feat] gather buffer (one per use: 'x' for forward, 'dy' for gradient)."""
        buf = self._full.get(key)
        if buf is None or buf.shape[1] != feat or buf.dtype != like.dtype or buf.device != like.device:
            buf = torch.empty((self.n, feat), dtype=like.dtype, device=like.device)
            self._full[key] = buf
        return buf

    def views(self, full):
        return [full[self.bounds[p]:self.bounds[p + 1]] for p in range(self.world)]

    def all_gather_rows(self, local, key):
        """[rows, feat] block of every rank -> [n, feat] on every rank (uneven block sizes)."""
        full = self.full_buffer(key, local.shape[1], local)
        views = self.views(full)
        if self.world == 1:
            if views[0].data_ptr() != local.data_ptr():
                views[0].copy_(local)
            return full
        if views[self.rank].data_ptr() != local.data_ptr():
            views[self.rank].copy_(local)
        # uneven blocks: one broadcast per owner (what NCCL's uneven all_gather does internally;
        # gloo has no uneven all_gather), issued together and awaited together
        works = [dist.broadcast(views[p], src=self._global_rank(p), group=self.group, async_op=True)
                 for p in range(self.world) if views[p].numel() > 0]
        for w in works:
            w.wait()
        return full

    def forward(self, x_local, out=None):
        """y_local = (A x)[r0:r1] for update_all(copy_src, sum)."""
        x_full = self.all_gather_rows(x_local, "x")
        return self._spmm(self.loc_in, x_full, out)

    def backward(self, dy_local, out=None):
        """dx_local = (A^T dy)[r0:r1]."""
        dy_full = self.all_gather_rows(dy_local, "dy")
        return self._spmm(self.loc_out, dy_full, out)
