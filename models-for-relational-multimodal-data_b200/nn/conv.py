"""Operator-level drop-ins for the torch_geometric classes the reference constructs:
``GINEConv``, ``PNAConv``, ``BatchNorm``, ``Linear`` (same constructor arguments, forward
signatures and state_dict keys), plus the edge-update step.

Reference call sites: /root/reference/src/nn/gnn/gine.py:18-19,66-77, src/nn/gnn/pna.py:23-30,
71-86, src/nn/models/{tabgnn.py:163-172, fused.py:204-214, inteleaved.py:189-199}.

Every conv offers ``forward(x, edge_index, edge_attr)`` (caller edge order, like PyG) and
``forward_sorted(x, seg, e)`` used by the backbones, which keep ``e`` in CSC storage order for
all layers so edge tensors stream sequentially.
"""
from __future__ import annotations

import math
from typing import List, Optional, Sequence, Union

import torch
import torch.nn.functional as F
from torch import Tensor, nn

from .. import functional as BF
from ..graph import EdgeStructure, SegmentIndex, edge_structure

Linear = nn.Linear     # torch_geometric.nn.Linear with known in_channels: same keys (weight, bias)

_AGG_ORDER = ("mean", "min", "max", "std")     # column order the kernel writes


class BatchNorm(nn.Module):
    """torch_geometric.nn.BatchNorm: BatchNorm1d under ``.module`` (keys ``module.*``)."""

    def __init__(self, in_channels: int, eps: float = 1e-5, momentum: float = 0.1, affine: bool = True,
                 track_running_stats: bool = True):
        super().__init__()
        self.module = nn.BatchNorm1d(in_channels, eps, momentum, affine, track_running_stats)

    def reset_parameters(self):
        self.module.reset_parameters()

    def forward(self, x: Tensor) -> Tensor:
        return self.module(x)


def _resolve(x, edge_index, reverse=False) -> EdgeStructure:
    xs = x[0] if isinstance(x, (tuple, list)) else x
    return edge_structure(edge_index, xs.shape[0], reverse=reverse)


class GINEConv(nn.Module):
    """out_i = nn((1+eps) x_i + sum_{j->i} relu(x_j + lin(e_ji)))  (PyG GINEConv)."""

    def __init__(self, nn: nn.Module, eps: float = 0.0, train_eps: bool = False, edge_dim: Optional[int] = None):
        super().__init__()
        self.nn = nn
        self.initial_eps = eps
        if train_eps:
            self.eps = torch.nn.Parameter(torch.empty(1))
        else:
            self.register_buffer("eps", torch.empty(1))
        first = nn[0] if isinstance(nn, torch.nn.Sequential) else nn
        in_channels = first.in_features
        if edge_dim is None:
            raise ValueError("b200mp GINEConv needs edge_dim (the reference always passes it: gine.py:18-19,66-70)")
        self.lin = Linear(edge_dim, in_channels)
        self.reset_parameters()

    def reset_parameters(self):
        for m in self.nn.modules():
            if hasattr(m, "reset_parameters") and m is not self.nn:
                m.reset_parameters()
        self.eps.data.fill_(self.initial_eps)
        self.lin.reset_parameters()

    def hoist(self, x: Tensor, gather):
        """the source-side node table of this conv on every rank's rows (multi-GPU): may be requested one step early, so that its
        all-gather runs behind the previous layer's edge update (parallel.AsyncRows)"""
        return gather(x)

    def forward_sorted(self, x: Union[Tensor, tuple], seg: SegmentIndex, e: Tensor, gather=None, sink=None, e_eid=None, hoisted=None) -> Tensor:
        """`gather` (multi-GPU, parallel.py): maps the owned rows of the source-side node table to
        the table of all ranks; the self term and the output stay on the owned rows.  `sink`:
        functional.EdgeGradSink shared with the layer's edge update.  `e_eid`: int32 row map when e
        is in the caller's edge order instead of segment order (functional.pna_aggregate)."""
        self_term = True
        if isinstance(x, (tuple, list)):
            x, x_dst = x
            self_term = x_dst is not None
        if self.lin.weight.shape[0] != self.lin.weight.shape[1] or x.shape[1] != self.lin.weight.shape[0]:
            raise ValueError("b200mp GINEConv kernels need edge_dim == in_channels")
        xs = hoisted if hoisted is not None else (gather(x) if gather is not None else x)
        if hasattr(xs, "ready"):
            xs = xs.ready()                   # an all-gather issued earlier: the consumer stream waits for it here
        out = BF.gine_aggregate(xs, e, self.lin.weight, self.lin.bias, seg, sink, e_eid)
        if self_term:
            out = torch.addcmul(out, x, 1 + self.eps)
        return self._apply_nn(out)

    def _apply_nn(self, out: Tensor) -> Tensor:
        """``self.nn`` (the reference builds Linear -> ReLU -> Linear, gine.py:66-70): its Linears run as node linears
        on the tcgen05 kernels (functional.linear_group; library GEMM for shapes those do not cover)"""
        if isinstance(self.nn, torch.nn.Sequential) and all(isinstance(m, (torch.nn.Linear, torch.nn.ReLU)) for m in self.nn):
            for m in self.nn:
                out = BF.linear_group(out, (m.weight,), (m.bias,))[0] if isinstance(m, torch.nn.Linear) else torch.relu(out)
            return out
        return self.nn(out)

    def forward(self, x, edge_index: Tensor, edge_attr: Tensor) -> Tensor:
        st = _resolve(x, edge_index)
        return self.forward_sorted(x, st.fwd, BF.gather_rows(edge_attr, st.fwd.perm))


class _DegreeScalerBuffers(nn.Module):
    """Holds ``avg_deg_lin`` / ``avg_deg_log`` under the PyG key prefix ``aggr_module.``."""

    def __init__(self, deg: Tensor):
        super().__init__()
        deg = deg.to(torch.float)
        total = float(deg.sum())
        bins = torch.arange(deg.numel(), device=deg.device)
        self.register_buffer("avg_deg_lin", torch.tensor(float((bins * deg).sum()) / total))
        self.register_buffer("avg_deg_log", torch.tensor(float(((bins + 1).log() * deg).sum()) / total))


class PNAConv(nn.Module):
    """PyG PNAConv for the only configuration the reference builds: towers=1, pre_layers=1,
    post_layers=1, divide_input=False, edge_dim == in_channels."""

    def __init__(self, in_channels: int, out_channels: int, aggregators: List[str], scalers: List[str], deg: Tensor,
                 edge_dim: Optional[int] = None, towers: int = 1, pre_layers: int = 1, post_layers: int = 1,
                 divide_input: bool = False, **kwargs):
        super().__init__()
        if towers != 1 or pre_layers != 1 or post_layers != 1 or divide_input:
            raise ValueError("b200mp PNAConv supports towers=1, pre_layers=1, post_layers=1, divide_input=False "
                             "(every constructor site of the reference)")
        if edge_dim is None or edge_dim != in_channels:
            raise ValueError("b200mp PNAConv needs edge_dim == in_channels (as at pna.py:71-74)")
        if deg is None:
            raise ValueError("PNAConv needs the in-degree histogram `deg`")
        for a in aggregators:
            if a not in _AGG_ORDER:
                raise ValueError(f"unsupported aggregator '{a}' (kernels provide {_AGG_ORDER})")
        for s in scalers:
            if s not in ("identity", "amplification", "attenuation", "linear", "inverse_linear"):
                raise ValueError(f"Unknown scaler '{s}'")
        self.in_channels, self.out_channels = in_channels, out_channels
        self.aggregators, self.scalers = list(aggregators), list(scalers)
        self.edge_dim, self.towers, self.divide_input = edge_dim, towers, divide_input
        self.F_in, self.F_out = in_channels, out_channels
        self.aggr_module = _DegreeScalerBuffers(deg)
        self.edge_encoder = Linear(edge_dim, self.F_in)
        self.pre_nns = nn.ModuleList([nn.Sequential(Linear(3 * self.F_in, self.F_in))])
        n_in = (len(aggregators) * len(scalers) + 1) * self.F_in
        self.post_nns = nn.ModuleList([nn.Sequential(Linear(n_in, self.F_out))])
        self.lin = Linear(out_channels, out_channels)

    def reset_parameters(self):
        self.edge_encoder.reset_parameters()
        self.pre_nns[0][0].reset_parameters()
        self.post_nns[0][0].reset_parameters()
        self.lin.reset_parameters()

    def hoist(self, x: Tensor, gather=None):
        """(A, B, out0): the node halves of pre_nn and the self term of post_nn on the N rows (SURVEY.md section 7 "hoisting"); with
        `gather` (multi-GPU) B covers every rank's rows -- requested one step early, its all-gather runs behind the previous layer's
        edge update (parallel.AsyncRows)"""
        Fi = self.F_in
        pre, post = self.pre_nns[0][0], self.post_nns[0][0]
        Wi, Wj = pre.weight[:, :Fi], pre.weight[:, Fi:2 * Fi]
        A, B, out = BF.linear_group(x, (Wi, Wj, post.weight[:, :Fi]), (None, None, post.bias))
        if gather is not None:            # multi-GPU: source rows of every rank (parallel.py)
            B = gather(B)
        return A, B, out

    def forward_sorted(self, x: Tensor, seg: SegmentIndex, e: Tensor, gather=None, sink=None, e_eid=None, hoisted=None) -> Tensor:
        Fi = self.F_in
        pre = self.pre_nns[0][0]
        Wc = pre.weight[:, 2 * Fi:]
        # pre_nn is one Linear without activation: hoist the node halves to N rows and fold the
        # edge encoder into a single H x H matrix (SURVEY.md section 7 "hoisting")
        post = self.post_nns[0][0]
        A, B, out = hoisted if hoisted is not None else self.hoist(x, gather)
        if hasattr(B, "ready"):
            B = B.ready()                     # an all-gather issued earlier: the consumer stream waits for it here
        Wp = Wc @ self.edge_encoder.weight
        bp = Wc @ self.edge_encoder.bias + pre.bias
        agg4, _ = BF.pna_aggregate(A, B, e, Wp, bp, seg, sink, e_eid)

        na = len(self.aggregators)
        cols = [_AGG_ORDER.index(a) for a in self.aggregators]
        lin_avg, log_avg = self.aggr_module.avg_deg_lin, self.aggr_module.avg_deg_log
        agg_used = agg4 if cols == [0, 1, 2, 3] else torch.cat([agg4[:, c * Fi:(c + 1) * Fi] for c in cols], dim=1)
        Ws = post.weight[:, Fi:]                          # [F_out, ns * na * Fi], scaler-major
        ns = len(self.scalers)
        # one GEMM for all scalers: Z[:, s*F_out:(s+1)*F_out] = agg_used @ Ws_s^T
        Z = BF.linear3x(agg_used, Ws.reshape(self.F_out, ns, na * Fi).permute(1, 0, 2).reshape(ns * self.F_out, na * Fi))
        out = BF.pna_scale(out, Z, seg.ptr, self.scalers, lin_avg, log_avg)
        return BF.linear_group(out, (self.lin.weight,), (self.lin.bias,))[0]

    def forward(self, x: Tensor, edge_index: Tensor, edge_attr: Tensor) -> Tensor:
        st = _resolve(x, edge_index)
        return self.forward_sorted(x, st.fwd, BF.gather_rows(edge_attr, st.fwd.perm))


def edge_update_sorted(emlp: nn.Sequential, x: Tensor, st: EdgeStructure, e: Tensor, alpha: float, beta: float,
                       gather=None, sink=None, e_eid=None, out_eid=None, hoisted=None) -> Tensor:
    """e' = alpha e + beta emlp([x[src], x[dst], e]) with emlp = Sequential(Linear(3H,H), ReLU,
    Linear(H,H)) (gine.py:71-75,88-89; pna.py:80-84,97-99; models/fused.py:215-219,254)."""
    l1, l2 = emlp[0], emlp[2]
    H = x.shape[1]
    Ws, Wd, We = l1.weight[:, :H], l1.weight[:, H:2 * H], l1.weight[:, 2 * H:]
    if hoisted is not None:               # (P on every rank's rows, Q): computed and gathered by the caller
        P, Q = hoisted
    else:
        P, Q = BF.linear_group(x, (Ws, Wd), (None, None))
        if gather is not None:
            P = gather(P)
    return BF.edge_update(P, Q, e, We, l1.bias, l2.weight, l2.bias, st, alpha, beta, sink, e_eid, out_eid)
