"""Multi-GPU message passing: destination-range partition with one exchange step per layer.

No counterpart in the reference (single GPU, no torch.distributed: SURVEY.md section 2.1); the
design is SURVEY.md section 8(e).  One process per GPU (torch.distributed, NCCL; gloo on CPU for
the host-logic tests).  Rank g owns a contiguous range of destination nodes balanced by in-edge
count, the rows of x for those nodes, its slice of the CSC structure and the features of every
edge whose destination it owns, for the whole backbone.  Per layer:

  forward   all-gather of the hoisted source-side node table (PNA: B = x Wj^T, edge update:
            P = x' Ws^T, GINE: x) -> every rank can gather any source row;
            all-reduce of the BatchNorm sums (batch statistics equal the single-GPU ones);
  backward  reduce-scatter of the gathered table's gradient (contributions to remote sources),
            all-reduce of the BatchNorm gradient sums, all-reduce of the weight gradients.

Parity rule: the G-rank result equals the 1-rank result up to fp32 re-association.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import torch
import torch.distributed as dist
import torch.nn.functional as F

from . import functional as BF
from . import graph as G
from .graph import EdgeStructure, SegmentIndex


@dataclass
class Partition:
    rank: int
    world: int
    num_nodes: int              # global N
    num_edges: int              # global E
    bounds: torch.Tensor        # int64 [world + 1] node range boundaries (host)
    n_max: int                  # largest owned range (padding unit of the gathered tables)
    st: EdgeStructure           # local structure: dst ids local, src ids in the padded gathered space
    edge_ids: torch.Tensor      # int64 [E_local] caller-order ids of the local edges, in local CSC order
    group: Optional[object] = None

    @property
    def lo(self) -> int:
        return int(self.bounds[self.rank])

    @property
    def hi(self) -> int:
        return int(self.bounds[self.rank + 1])

    @property
    def n_local(self) -> int:
        return self.hi - self.lo

    def shard_nodes(self, x: torch.Tensor) -> torch.Tensor:
        return x[self.lo:self.hi].contiguous()

    def shard_edges(self, e: torch.Tensor) -> torch.Tensor:
        """Rows of a caller-order edge tensor owned by this rank, in local CSC storage order."""
        return e[self.edge_ids].contiguous()


def partition_graph(edge_index: torch.Tensor, num_nodes: int, rank: int, world: int, group=None) -> Partition:
    """Split the destination nodes into `world` contiguous ranges with ~E/world in-edges each."""
    full = G.build_segment_index(edge_index[1], edge_index[0], num_nodes)
    E = int(full.num_edges)
    targets = torch.arange(1, world, device=full.ptr.device, dtype=torch.int64) * E // world
    inner = torch.searchsorted(full.ptr, targets, right=False).clamp_(max=num_nodes)
    bounds = torch.cat([torch.zeros(1, dtype=torch.int64), inner.cpu(), torch.tensor([num_nodes], dtype=torch.int64)])
    bounds = torch.cummax(bounds, 0).values
    sizes = bounds[1:] - bounds[:-1]
    n_max = max(int(sizes.max()), 1)
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    e0, e1 = int(full.ptr[lo]), int(full.ptr[hi])
    dev = full.ptr.device
    # source ids -> index in the padded all-gathered table: owner * n_max + (src - lo_owner)
    src = full.other[e0:e1]
    b = bounds.to(dev)
    owner = torch.searchsorted(b, src, right=True) - 1
    src_pad = owner * n_max + (src - b[owner])
    ptr = (full.ptr[lo:hi + 1] - e0).contiguous()
    seg32 = (full.seg32[e0:e1] - lo).contiguous()
    fwd = SegmentIndex(hi - lo, e1 - e0, torch.arange(e1 - e0, device=dev), ptr, src_pad.contiguous(),
                       src_pad.to(torch.int32).contiguous(), seg32)
    st = EdgeStructure(hi - lo, e1 - e0, fwd)
    return Partition(rank, world, int(num_nodes), E, bounds, n_max, st, full.perm[e0:e1].contiguous(), group)


# ----------------------------------------------------------------------------- copy-engine all-gather
_CE = {"state": None, "group": None}


def _ce_gather_ready(part: "Partition") -> bool:
    """Copy-engine all-gather over peer memory (torch's symmetric-memory `_low_contention_all_gather`: every rank publishes its shard
    in a P2P-mapped workspace, the others PULL it with copy-engine memcpys on a side stream, barriers on the signal pads): no SMs are
    taken from the persistent edge kernels, which is what made the NCCL all-gather impossible to overlap (8 x B200: the edge-update
    forward slowed from 1.22 to 1.58 ms when the NCCL gather ran beside it).  Opt-in with B200MP_CE_GATHER=1, default process group,
    NCCL backend only; any failure at set-up falls back to the NCCL collective for the rest of the run."""
    import os
    if _CE["state"] is None:
        ok = False
        if os.environ.get("B200MP_CE_GATHER", "0") == "1" and part.group is None and dist.get_backend() == "nccl":
            try:
                import torch.distributed._symmetric_memory as symm
                name = dist.group.WORLD.group_name
                symm.enable_symm_mem_for_group(name)
                _CE["group"] = name
                ok = True
            except Exception as ex:          # pragma: no cover (needs a multi-GPU box)
                import sys
                print(f"[b200mp] copy-engine all-gather unavailable ({type(ex).__name__}: {ex}); using NCCL", file=sys.stderr)
        _CE["state"] = ok
    return bool(_CE["state"])


def _ce_all_gather(mine: torch.Tensor):
    """-> gathered [world * n_max, H] still in flight on the symmetric-memory stream (wait_tensor before use), or None on failure"""
    try:
        return torch.ops.symm_mem._low_contention_all_gather(mine, _CE["group"])
    except Exception as ex:                  # pragma: no cover
        import sys
        print(f"[b200mp] copy-engine all-gather failed ({type(ex).__name__}: {ex}); using NCCL from here on", file=sys.stderr)
        _CE["state"] = False
        return None


# ----------------------------------------------------------------------------- collectives with autograd
class _AllGatherRows(torch.autograd.Function):
    """[n_local, H] -> padded [world * n_max, H]; backward = reduce-scatter of the gradient."""

    @staticmethod
    def forward(ctx, x, part: Partition):
        ctx.part = part
        H = x.shape[1]
        mine = x.new_zeros((part.n_max, H))
        mine[: x.shape[0]] = x
        if _ce_gather_ready(part):
            out = _ce_all_gather(mine)
            if out is not None:
                return torch.ops._c10d_functional.wait_tensor(out)
        out = x.new_empty((part.world * part.n_max, H))
        dist.all_gather_into_tensor(out, mine, group=part.group)
        return out

    @staticmethod
    def backward(ctx, g):
        part: Partition = ctx.part
        g = g.contiguous()
        if dist.get_backend(part.group) == "nccl":
            out = g.new_empty((part.n_max, g.shape[1]))
            dist.reduce_scatter_tensor(out, g, group=part.group)
        else:                                   # gloo has no reduce-scatter: all-reduce + slice
            dist.all_reduce(g, group=part.group)
            out = g[part.rank * part.n_max:(part.rank + 1) * part.n_max]
        return out[: part.n_local].contiguous(), None


def all_gather_rows(x, part: Partition):
    return _AllGatherRows.apply(x, part)


class _CEWork:
    """wait() of a copy-engine gather: the current stream waits for the symmetric-memory stream"""

    def __init__(self, out):
        self.out = out

    def wait(self):
        torch.ops._c10d_functional.wait_tensor(self.out)


class AsyncRows:
    """An all-gather of node rows in flight on the communication stream.  ``ready()`` makes the current stream wait for it and
    returns the gathered table (autograd-connected: its backward is the reduce-scatter of ``_AllGatherRows``)."""

    def __init__(self, table, work, keep=None):
        self.table, self.work, self.keep = table, work, keep          # keep: the send buffer, alive until the collective is done

    def ready(self):
        if self.work is not None:
            self.work.wait()
            self.work = self.keep = None
        return self.table


class _AllGatherRowsAsync(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, part: Partition, box):
        ctx.part = part
        H = x.shape[1]
        mine = x.new_zeros((part.n_max, H))
        mine[: x.shape[0]] = x
        if _ce_gather_ready(part):
            out = _ce_all_gather(mine)
            if out is not None:
                box.append(_CEWork(out))
                box.append(mine)
                return out
        out = x.new_empty((part.world * part.n_max, H))
        box.append(dist.all_gather_into_tensor(out, mine, group=part.group, async_op=True))
        box.append(mine)                       # keep the send buffer alive until the collective is done
        return out

    @staticmethod
    def backward(ctx, g):
        return _AllGatherRows.backward(ctx, g) + (None,)


def all_gather_rows_async(x, part: Partition) -> AsyncRows:
    """start the all-gather now, use it later: between the two the current stream keeps computing (the gather of layer l + 1's
    source table runs behind layer l's edge-update kernel)"""
    box = []
    table = _AllGatherRowsAsync.apply(x, part, box)
    return AsyncRows(table, box[0], box[1])


class _SyncBatchNorm(torch.autograd.Function):
    """BatchNorm1d training forward over the rows of ALL ranks (sums all-reduced)."""

    @staticmethod
    def forward(ctx, x, weight, bias, eps, part: Partition):
        stats = torch.stack([x.sum(0), (x * x).sum(0)])
        dist.all_reduce(stats, group=part.group)
        n = float(part.num_nodes)
        mean = stats[0] / n
        var = (stats[1] / n - mean * mean).clamp_(min=0.0)
        invstd = torch.rsqrt(var + eps)
        xhat = (x - mean) * invstd
        ctx.save_for_backward(xhat, invstd, weight)
        ctx.part = part
        ctx.mark_non_differentiable(mean, var)
        return xhat * weight + bias, mean, var

    @staticmethod
    def backward(ctx, g, _gm, _gv):
        xhat, invstd, weight = ctx.saved_tensors
        part: Partition = ctx.part
        sums = torch.stack([g.sum(0), (g * xhat).sum(0)])
        gw, gb = sums[1].clone(), sums[0].clone()          # local parts; all-reduced with the other weight grads
        dist.all_reduce(sums, group=part.group)
        n = float(part.num_nodes)
        gx = (g - sums[0] / n - xhat * (sums[1] / n)) * (invstd * weight)
        return gx, gw, gb, None, None


def sync_batch_norm(bn: torch.nn.BatchNorm1d, x, part: Partition):
    if not bn.training:
        return F.batch_norm(x, bn.running_mean, bn.running_var, bn.weight, bn.bias, False, 0.0, bn.eps)
    y, mean, var = _SyncBatchNorm.apply(x, bn.weight, bn.bias, bn.eps, part)
    if bn.track_running_stats:
        with torch.no_grad():
            n = float(part.num_nodes)
            m = bn.momentum if bn.momentum is not None else 0.1
            bn.running_mean.mul_(1 - m).add_(mean, alpha=m)
            bn.running_var.mul_(1 - m).add_(var * (n / max(n - 1.0, 1.0)), alpha=m)
            bn.num_batches_tracked += 1
    return y


def all_reduce_gradients(module: torch.nn.Module, part: Partition):
    """Sum the parameter gradients over the ranks (each holds the contribution of its edges/nodes)."""
    grads = [p.grad for p in module.parameters() if p.grad is not None]
    if not grads:
        return
    flat = torch.cat([g.reshape(-1) for g in grads])
    dist.all_reduce(flat, group=part.group)
    o = 0
    for g in grads:
        g.copy_(flat[o:o + g.numel()].view_as(g))
        o += g.numel()
