"""ORACLE (test infrastructure, not product code).

CPU restatement, in plain PyTorch, of the third-party operators the reference's hot
path calls: torch_geometric 2.5.3 ``GINEConv``, ``PNAConv`` (+ ``DegreeScalerAggregation``,
Mean/Min/Max/Std aggregations), ``BatchNorm``, ``Linear`` and ``utils.degree``.

torch_geometric is NOT vendored under /root/reference (environment.yml:336 pins
``torch-geometric==2.5.3``) and is not installable here, so its published algorithm is
restated from upstream semantics.  **Parity unpinned** at this
boundary: the reference holds no golden vectors or asserting tests for these ops
(SURVEY.md §4, §8c).  What IS pinned: the reference's own model files
(src/nn/gnn/gine.py, src/nn/gnn/pna.py, src/nn/models/*.py) run unchanged on top of these
functions through ``oracle/shim`` and produce the fixtures in ``tests/golden``
(see ``oracle/gen_golden.py``); ``oracle/models.py`` is checked against those fixtures.

Every function is dtype-generic (fp32 for parity, fp64 as the round-off yardstick) and
written functionally over a ``{state_dict key: tensor}`` mapping so that the very same
state_dict drives the reference modules, this oracle and the CUDA drop-ins.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg
may import this package.
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Sequence

import torch
import torch.nn.functional as F
from torch import Tensor

Params = Dict[str, Tensor]


def sub(p: Params, prefix: str) -> Params:
    """View of a state_dict restricted to ``prefix`` (prefix stripped)."""
    n = len(prefix)
    return {k[n:]: v for k, v in p.items() if k.startswith(prefix)}


# --------------------------------------------------------------------------- utils
def degree(index: Tensor, num_nodes: int, dtype=torch.float32) -> Tensor:
    """torch_geometric.utils.degree: scatter_add of ones (called at main.py:286 and inside
    DegreeScalerAggregation.forward)."""
    out = torch.zeros(num_nodes, dtype=dtype, device=index.device)
    return out.scatter_add_(0, index, torch.ones(index.numel(), dtype=dtype, device=index.device))


def _bcast(index: Tensor, like: Tensor) -> Tensor:
    return index.view(-1, *([1] * (like.dim() - 1))).expand_as(like)


def scatter_sum(src: Tensor, index: Tensor, dim_size: int) -> Tensor:
    out = src.new_zeros((dim_size,) + tuple(src.shape[1:]))
    return out.scatter_add_(0, _bcast(index, src), src)


def scatter_mean(src: Tensor, index: Tensor, dim_size: int) -> Tensor:
    """torch_geometric.utils.scatter(reduce='mean'): sum / clamp(count, 1)."""
    count = degree(index, dim_size, src.dtype).clamp_(min=1)
    out = scatter_sum(src, index, dim_size)
    return out / count.view(-1, *([1] * (src.dim() - 1)))


def scatter_minmax(src: Tensor, index: Tensor, dim_size: int, which: str) -> Tensor:
    """torch_geometric.utils.scatter(reduce='min'|'max') without torch_scatter:
    zeros.scatter_reduce_(amin|amax, include_self=False) -> empty segments stay 0 and the
    gradient is split evenly between ties."""
    out = src.new_zeros((dim_size,) + tuple(src.shape[1:]))
    return out.scatter_reduce_(0, _bcast(index, src), src, reduce="a" + which, include_self=False)


def aggr_std(src: Tensor, index: Tensor, dim_size: int) -> Tensor:
    """StdAggregation(VarAggregation(semi_grad=False)): sqrt(clamp(E[m^2]-E[m]^2, 1e-5)),
    then masked to 0 where <= sqrt(1e-5)."""
    mean = scatter_mean(src, index, dim_size)
    mean2 = scatter_mean(src * src, index, dim_size)
    var = mean2 - mean * mean
    out = var.clamp(min=1e-5).sqrt()
    return out.masked_fill(out <= math.sqrt(1e-5), 0.0)


_AGGRS = {
    "mean": scatter_mean,
    "sum": scatter_sum,
    "add": scatter_sum,
    "min": lambda s, i, n: scatter_minmax(s, i, n, "min"),
    "max": lambda s, i, n: scatter_minmax(s, i, n, "max"),
    "std": aggr_std,
}


def avg_deg_from_hist(deg_hist: Tensor):
    """DegreeScalerAggregation.__init__: (avg_deg_lin, avg_deg_log) from the training-graph
    in-degree histogram built at utils.py:179-182."""
    d = deg_hist.to(torch.float)
    total = int(d.sum())
    bins = torch.arange(d.numel(), device=d.device)
    lin = float((bins * d).sum()) / total
    log = float(((bins + 1).log() * d).sum()) / total
    return lin, log


def degree_scaler_aggregation(msg: Tensor, index: Tensor, dim_size: int,
                              aggregators: Sequence[str], scalers: Sequence[str],
                              avg_deg_lin: Tensor, avg_deg_log: Tensor) -> Tensor:
    """DegreeScalerAggregation.forward (MultiAggregation mode='cat', then scalers, cat)."""
    out = torch.cat([_AGGRS[a](msg, index, dim_size) for a in aggregators], dim=-1)
    deg = degree(index, dim_size, out.dtype).view(-1, *([1] * (out.dim() - 1)))
    outs = []
    for s in scalers:
        if s == "identity":
            o = out
        elif s == "amplification":
            o = out * (torch.log(deg + 1) / avg_deg_log)
        elif s == "attenuation":
            o = out * (avg_deg_log / torch.log(deg.clamp(min=1) + 1))
        elif s == "linear":
            o = out * (deg / avg_deg_lin)
        elif s == "inverse_linear":
            o = out * (avg_deg_lin / deg.clamp(min=1))
        else:
            raise ValueError(f"Unknown scaler '{s}'")
        outs.append(o)
    return torch.cat(outs, dim=-1) if len(outs) > 1 else outs[0]


# --------------------------------------------------------------------------- convs
def gine_conv(p: Params, x: Tensor, edge_index: Tensor, edge_attr: Tensor,
              self_term: bool = True) -> Tensor:
    """GINEConv.forward with nn = Sequential(Linear, ReLU, Linear) and edge_dim set, as built
    at src/nn/gnn/gine.py:60-70.  ``self_term=False`` is the ``(x, None)`` call at
    gine.py:32-33 (no (1+eps)*x_i term).  Keys: nn.0.*, nn.2.*, lin.*, eps."""
    src, dst = edge_index[0], edge_index[1]
    msg = (x[src] + F.linear(edge_attr, p["lin.weight"], p["lin.bias"])).relu()
    out = scatter_sum(msg, dst, x.shape[0])
    if self_term:
        out = out + (1 + p["eps"].to(x.dtype)) * x
    h = F.linear(out, p["nn.0.weight"], p["nn.0.bias"]).relu()
    return F.linear(h, p["nn.2.weight"], p["nn.2.bias"])


def pna_conv(p: Params, x: Tensor, edge_index: Tensor, edge_attr: Tensor,
             aggregators: Sequence[str], scalers: Sequence[str]) -> Tensor:
    """PNAConv.forward for towers=1, pre_layers=1, post_layers=1, divide_input=False (the only
    configuration the reference constructs: src/nn/gnn/pna.py:71-79, models/tabgnn.py:163-172,
    models/fused.py:204-214, models/inteleaved.py:189-199).
    Keys: edge_encoder.*, pre_nns.0.0.*, post_nns.0.0.*, lin.*, aggr_module.avg_deg_{lin,log}."""
    src, dst = edge_index[0], edge_index[1]
    n = x.shape[0]
    x3 = x.view(n, 1, -1)
    e = F.linear(edge_attr, p["edge_encoder.weight"], p["edge_encoder.bias"]).view(-1, 1, x.shape[1])
    h = torch.cat([x3[dst], x3[src], e], dim=-1)               # [x_i, x_j, e]
    msg = F.linear(h, p["pre_nns.0.0.weight"], p["pre_nns.0.0.bias"])   # [E,1,F]
    agg = degree_scaler_aggregation(msg, dst, n, aggregators, scalers,
                                    p["aggr_module.avg_deg_lin"].to(x.dtype),
                                    p["aggr_module.avg_deg_log"].to(x.dtype))
    out = torch.cat([x3, agg], dim=-1)[:, 0]
    out = F.linear(out, p["post_nns.0.0.weight"], p["post_nns.0.0.bias"])
    return F.linear(out, p["lin.weight"], p["lin.bias"])


def batch_norm(p: Params, x: Tensor, training: bool, momentum: float = 0.1,
               eps: float = 1e-5, prefix: str = "module.") -> Tensor:
    """torch_geometric.nn.BatchNorm -> torch.nn.BatchNorm1d (keys module.*).  Updates the
    running stats held in ``p`` in place when training, like the module does."""
    rm, rv = p.get(prefix + "running_mean"), p.get(prefix + "running_var")
    if rm is not None and rm.dtype != x.dtype:       # fp64 yardstick runs: keep stats local
        rm, rv = rm.to(x.dtype), rv.to(x.dtype)
    out = F.batch_norm(x, rm, rv, p[prefix + "weight"], p[prefix + "bias"],
                       training, momentum, eps)
    if training and prefix + "num_batches_tracked" in p:
        p[prefix + "num_batches_tracked"] += 1
    return out


def hetero(conv_fn, p: Params, x: Tensor, edge_index: Tensor, edge_attr: Tensor, **kw) -> Tensor:
    """GINEConvHetero / PNAConvHetero (gine.py:15-35, pna.py:17-46): forward conv on
    edge_index, backward conv on edge_index.flipud(), Linear(3H->H) on [x, a_in, a_out]."""
    a_in = conv_fn(sub(p, "conv_forw."), x, edge_index, edge_attr, **kw)
    a_out = conv_fn(sub(p, "conv_back."), x, edge_index.flipud(), edge_attr, **kw)
    return F.linear(torch.cat([x, a_in, a_out], dim=1), p["lin.weight"], p["lin.bias"])
