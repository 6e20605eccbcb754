import torch
from oracle import pyg_functional as PF
from .dense.linear import Linear
from .inits import reset
from . import conv, dense, inits  # noqa: F401


class BatchNorm(torch.nn.Module):
    """torch_geometric.nn.BatchNorm: wraps BatchNorm1d as ``self.module``."""

    def __init__(self, in_channels, eps=1e-5, momentum=0.1, affine=True,
                 track_running_stats=True, allow_single_element=False):
        super().__init__()
        self.module = torch.nn.BatchNorm1d(in_channels, eps, momentum, affine, track_running_stats)
        self.in_channels = in_channels

    def reset_parameters(self):
        self.module.reset_parameters()

    def forward(self, x):
        return self.module(x)


class GINEConv(torch.nn.Module):
    def __init__(self, nn, eps=0.0, train_eps=False, edge_dim=None, **kw):
        super().__init__()
        self.nn = nn
        self.initial_eps = eps
        if train_eps:
            self.eps = torch.nn.Parameter(torch.empty(1))
        else:
            self.register_buffer("eps", torch.empty(1))
        first = nn[0] if isinstance(nn, torch.nn.Sequential) else nn
        self.lin = Linear(edge_dim, first.in_features) if edge_dim is not None else None
        self.reset_parameters()

    def reset_parameters(self):
        reset(self.nn)
        self.eps.data.fill_(self.initial_eps)
        if self.lin is not None:
            self.lin.reset_parameters()

    def forward(self, x, edge_index, edge_attr=None, size=None):
        self_term = True
        if isinstance(x, (tuple, list)):
            self_term = x[1] is not None
            x = x[0]
        p = dict(self.named_parameters())
        p.update(dict(self.named_buffers()))
        return PF.gine_conv(p, x, edge_index, edge_attr, self_term=self_term)


class _AggrHolder(torch.nn.Module):
    def __init__(self, deg):
        super().__init__()
        lin, log = PF.avg_deg_from_hist(deg)
        self.register_buffer("avg_deg_lin", torch.tensor([lin]))
        self.register_buffer("avg_deg_log", torch.tensor([log]))


class PNAConv(torch.nn.Module):
    def __init__(self, in_channels, out_channels, aggregators, scalers, deg, edge_dim=None,
                 towers=1, pre_layers=1, post_layers=1, divide_input=False, **kw):
        super().__init__()
        assert towers == 1 and pre_layers == 1 and post_layers == 1 and not divide_input
        assert in_channels == out_channels and edge_dim is not None
        self.aggregators, self.scalers = list(aggregators), list(scalers)
        self.aggr_module = _AggrHolder(deg)
        F_in = in_channels
        self.edge_encoder = Linear(edge_dim, F_in)
        self.pre_nns = torch.nn.ModuleList([torch.nn.Sequential(Linear(3 * F_in, F_in))])
        self.post_nns = torch.nn.ModuleList([torch.nn.Sequential(
            Linear((len(aggregators) * len(scalers) + 1) * F_in, out_channels))])
        self.lin = Linear(out_channels, out_channels)

    def reset_parameters(self):
        self.edge_encoder.reset_parameters()
        for nn in self.pre_nns:
            reset(nn)
        for nn in self.post_nns:
            reset(nn)
        self.lin.reset_parameters()

    def forward(self, x, edge_index, edge_attr=None):
        p = dict(self.named_parameters())
        p.update(dict(self.named_buffers()))
        return PF.pna_conv(p, x, edge_index, edge_attr, self.aggregators, self.scalers)
