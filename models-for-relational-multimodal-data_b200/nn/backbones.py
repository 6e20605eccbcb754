"""Drop-in backbones: ``GINe`` (/root/reference/src/nn/gnn/gine.py:37-100), ``PNAS``
(src/nn/gnn/pna.py:48-101) and their reverse-message-passing convs ``GINEConvHetero``
(gine.py:15-35) / ``PNAConvHetero`` (pna.py:17-46).  Same constructors, ``forward(x, edge_index,
edge_attr) -> (x, edge_attr)`` and state_dict keys.

The edge tensor is permuted once into CSC storage order, stays there for all layers and is
permuted back at the end; all per-edge work runs in the b200mp CUDA kernels.
"""
from __future__ import annotations

import os as _os

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import functional as BF
from ..graph import EdgeStructure, edge_structure
from .conv import BatchNorm, GINEConv, Linear, PNAConv, edge_update_sorted


class _HeteroMixin:
    def reset_parameters(self):
        self.conv_forw.reset_parameters()
        self.conv_back.reset_parameters()
        self.lin.reset_parameters()


class GINEConvHetero(_HeteroMixin, nn.Module):
    def __init__(self, network, n_hidden):
        super().__init__()
        self.conv_forw = GINEConv(network, edge_dim=n_hidden)
        self.conv_back = GINEConv(network, edge_dim=n_hidden)      # same `network` object, as in the reference
        self.lin = Linear(n_hidden * 3, n_hidden)

    def forward_sorted(self, x, st: EdgeStructure, e, sink=None, e_eid=None, rev_eid=None):
        """e_eid / rev_eid: row maps of the two convs when e is in the caller's edge order (first layer of the backbones)"""
        a_in = self.conv_forw.forward_sorted((x, None), st.fwd, e, sink=sink, e_eid=e_eid)
        a_out = self.conv_back.forward_sorted((x, None), st.rev, e, e_eid=rev_eid)
        return BF.linear_cat([x, a_in, a_out], self.lin.weight, self.lin.bias)

    def forward(self, x, edge_index, edge_attr):
        st = edge_structure(edge_index, x.shape[0], reverse=True)
        return self.forward_sorted(x, st, BF.gather_rows(edge_attr, st.fwd.perm))


class PNAConvHetero(_HeteroMixin, nn.Module):
    def __init__(self, n_hidden, in_channels, out_channels, aggregators, scalers, deg, edge_dim, towers=1,
                 pre_layers=1, post_layers=1, divide_input=False):
        super().__init__()
        kw = dict(in_channels=in_channels, out_channels=out_channels, aggregators=aggregators, scalers=scalers,
                  deg=deg, edge_dim=edge_dim, towers=towers, pre_layers=pre_layers, post_layers=post_layers,
                  divide_input=divide_input)
        self.conv_forw = PNAConv(**kw)
        self.conv_back = PNAConv(**kw)
        self.lin = Linear(n_hidden * 3, n_hidden)

    def forward_sorted(self, x, st: EdgeStructure, e, sink=None, e_eid=None, rev_eid=None):
        a_in = self.conv_forw.forward_sorted(x, st.fwd, e, sink=sink, e_eid=e_eid)
        a_out = self.conv_back.forward_sorted(x, st.rev, e, e_eid=rev_eid)
        return BF.linear_cat([x, a_in, a_out], self.lin.weight, self.lin.bias)

    def forward(self, x, edge_index, edge_attr):
        st = edge_structure(edge_index, x.shape[0], reverse=True)
        return self.forward_sorted(x, st, BF.gather_rows(edge_attr, st.fwd.perm))


def _conv_sorted(conv, x, st: EdgeStructure, e, sink=None):
    if isinstance(conv, (GINEConvHetero, PNAConvHetero)):
        return conv.forward_sorted(x, st, e, sink=sink)
    return conv.forward_sorted(x, st.fwd, e, sink=sink)


class _EdgeStorage:
    """bf16 STORAGE of the [E,H] edge tensors inside the layer loop (BASELINE.json config 3): ``model.set_edge_storage(torch.bfloat16)``.
    Node tensors, parameters and all arithmetic stay fp32; the embedded edge features are rounded to bf16 once, every edge tensor
    produced inside (messages, hidden rows, updated features, their gradients) is stored as bf16, and the model returns the updated
    edge features in that dtype.  H = 64 (tensor-core kernels) only."""
    edge_dtype = torch.float32

    def set_edge_storage(self, dtype):
        if dtype not in (torch.float32, torch.bfloat16):
            raise ValueError("edge storage is torch.float32 or torch.bfloat16")
        self.edge_dtype = dtype
        return self

    def _store(self, e):
        return e if e.dtype == self.edge_dtype else e.to(self.edge_dtype)


def _embed(lin: nn.Linear, t):
    """node_emb / edge_emb (gine.py:83-84, pna.py:91-93): 64 -> 64 embeddings of >= 4096 rows run on the tcgen05 node kernel
    (3xTF32, one pass over HBM; functional.linear_group falls back to the library GEMM for other shapes)"""
    return BF.linear_group(t, (lin.weight,), (lin.bias,))[0]


def _embed_wide(lin: nn.Linear, t, scale: float = 1.0):
    """lin(scale * t) for the flattened column tokens [rows, S C] (tabgnn.py:130-131 and its twins: `edge_emb(((e + te) / 2).reshape(..))`):
    the scale (a power of two, exact) is folded into the tiny weight instead of costing a pass over the [rows, S C] tensor.  The product
    itself stays on the library's fp32 GEMM: measured on the TABGNN C = H = 64 parity case, the 3xTF32 node kernels here (~1e-6 per
    product against fp32's 6e-8) are amplified ~100x by the last BatchNorm's weight-gradient sum (2e-4 against the oracle's 7e-6, every
    activation still at 2e-6), for 11 ms of a 471 ms step.  B200MP_EMB_TC=1 selects the tcgen05 node kernels."""
    W = lin.weight if scale == 1.0 else lin.weight * scale
    if _os.environ.get("B200MP_EMB_TC", "0") != "1":
        return torch.nn.functional.linear(t, W, lin.bias)
    if t.shape[1] == 64:
        return BF.linear_group(t, (W,), (lin.bias,))[0]
    return BF.linear3x(t, W, lin.bias)


class _Backbone(_EdgeStorage, nn.Module):
    """Shared layer loop of gine.py:86-89 / pna.py:95-99."""

    def _layers(self, x, edge_index, e):
        st = edge_structure(edge_index, x.shape[0], reverse=self.reverse_mp)
        H = x.shape[1]
        e = self._store(e)
        if not (H == 64 or BF.wide_tc_available(H)) or not BF.row_maps_available():
            return self._layers_permuted(x, st, e)
        # e stays in the caller's edge order at both ends: the first layer's kernels read it through
        # the CSC permutation, the last edge update writes through it -- no permutation passes
        perm32 = st.fwd.perm32()
        rperm32 = st.rev.perm32() if self.reverse_mp else None          # reverse convs: CSR position -> caller row
        L = self.num_gnn_layers
        e_in = e
        for i in range(L):
            e_eid = perm32 if (i == 0 or not self.edge_updates) else None
            out_eid = perm32 if i == L - 1 else None
            sink = BF.EdgeGradSink() if (self.edge_updates and e.requires_grad and torch.is_grad_enabled()) else None
            if self.reverse_mp:
                c = self.convs[i].forward_sorted(x, st, e, sink=sink, e_eid=e_eid, rev_eid=rperm32 if e_eid is not None else None)
            else:
                c = self.convs[i].forward_sorted(x, st.fwd, e, sink=sink, e_eid=e_eid)
            x = BF.bn_relu_residual(self.batch_norms[i].module, c, x)
            if self.edge_updates:
                e = edge_update_sorted(self.emlps[i], x, st, e, 1.0, 0.5, sink=sink, e_eid=e_eid, out_eid=out_eid)
        return x, (e if self.edge_updates else e_in)

    def _layers_permuted(self, x, st, e):
        e = BF.gather_rows(e, st.fwd.perm)                      # caller order -> CSC storage order
        for i in range(self.num_gnn_layers):
            # conv and edge update of a layer read the same e: one gradient buffer for both
            sink = BF.EdgeGradSink() if (self.edge_updates and e.requires_grad and torch.is_grad_enabled()) else None
            x = BF.bn_relu_residual(self.batch_norms[i].module, _conv_sorted(self.convs[i], x, st, e, sink), x)
            if self.edge_updates:
                e = edge_update_sorted(self.emlps[i], x, st, e, 1.0, 0.5, sink=sink)
        return x, BF.scatter_rows(e, st.fwd.perm)               # back to caller order

    def message_passing_partitioned(self, x_local, e_local, part):
        """The same loop on one rank of a destination-range partition (parallel.Partition):
        x_local = owned node rows, e_local = owned edges in local CSC order (part.shard_*)."""
        from .. import parallel as PL
        if self.reverse_mp:
            raise NotImplementedError("reverse_mp is not supported on a partitioned graph")
        st, x, e = part.st, x_local, self._store(e_local)
        gather = lambda t: PL.all_gather_rows(t, part)
        gather_async = lambda t: PL.all_gather_rows_async(t, part)
        pre = None
        L = self.num_gnn_layers
        for i in range(L):
            sink = BF.EdgeGradSink() if (self.edge_updates and e.requires_grad and torch.is_grad_enabled()) else None
            c = self.convs[i].forward_sorted(x, st.fwd, e, gather=gather, sink=sink, hoisted=pre)
            pre = None
            bn = self.batch_norms[i].module
            if bn.training:
                x = BF.bn_relu_residual(bn, c, x, group=part.group, n_total=part.num_nodes)
            else:
                x = (x + F.relu(PL.sync_batch_norm(bn, c, part))) / 2
            if self.edge_updates:
                # the update needs ITS source table now; the next conv's table is requested right behind it, so that its all-gather
                # runs on the communication stream while the edge-update kernel computes
                l1 = self.emlps[i][0]
                H = x.shape[1]
                P, Q = BF.linear_group(x, (l1.weight[:, :H], l1.weight[:, H:2 * H]), (None, None))
                P = gather(P)
                if i + 1 < L:
                    pre = self.convs[i + 1].hoist(x, gather_async)
                e = edge_update_sorted(self.emlps[i], x, st, e, 1.0, 0.5, sink=sink, hoisted=(P, Q))
        return x, e

    def message_passing(self, x, edge_index, e):
        """The L-layer loop on already embedded x[N,H], e[E,H] (the region BASELINE.json's metric
        times: SURVEY.md section 8d)."""
        return self._layers(x, edge_index, e)


class GINe(_Backbone):
    def __init__(self, num_features=1, num_gnn_layers=2, n_hidden=100, edge_updates=False, edge_dim=None,
                 reverse_mp=False):
        super().__init__()
        self.n_hidden = n_hidden
        self.num_gnn_layers = num_gnn_layers
        self.edge_updates = edge_updates
        self.reverse_mp = reverse_mp

        self.node_emb = nn.Linear(num_features, n_hidden)
        self.edge_emb = nn.Linear(edge_dim, n_hidden)

        self.convs = nn.ModuleList()
        self.emlps = nn.ModuleList()
        self.batch_norms = nn.ModuleList()
        for _ in range(self.num_gnn_layers):
            net = nn.Sequential(nn.Linear(n_hidden, n_hidden), nn.ReLU(), nn.Linear(n_hidden, n_hidden))
            conv = GINEConvHetero(net, n_hidden=n_hidden) if reverse_mp else GINEConv(net, edge_dim=n_hidden)
            if self.edge_updates:
                self.emlps.append(nn.Sequential(nn.Linear(3 * n_hidden, n_hidden), nn.ReLU(),
                                                nn.Linear(n_hidden, n_hidden)))
            self.convs.append(conv)
            self.batch_norms.append(BatchNorm(n_hidden))

    def forward(self, x, edge_index, edge_attr):
        x = _embed(self.node_emb, x)
        e = _embed(self.edge_emb, edge_attr.view(edge_attr.shape[0], -1))
        return self._layers(x, edge_index, e)

    def reset_parameters(self):
        for conv in self.convs:
            conv.reset_parameters()
        for emlp in self.emlps:
            for m in emlp:
                if hasattr(m, "reset_parameters"):
                    m.reset_parameters()
        for bn in self.batch_norms:
            bn.reset_parameters()
        self.node_emb.reset_parameters()
        self.edge_emb.reset_parameters()


class PNAS(_Backbone):
    def __init__(self, num_features, num_gnn_layers, n_classes=2, n_hidden=128, edge_updates=True, edge_dim=None,
                 dropout=0.0, final_dropout=0.5, deg=None, reverse_mp=False):
        super().__init__()
        self.n_hidden = n_hidden
        self.num_gnn_layers = num_gnn_layers
        self.edge_updates = edge_updates
        self.final_dropout = final_dropout
        self.reverse_mp = reverse_mp

        aggregators = ['mean', 'min', 'max', 'std']
        scalers = ['identity', 'amplification', 'attenuation']

        self.node_emb = nn.Linear(num_features, n_hidden)
        self.edge_emb = nn.Linear(edge_dim, n_hidden)

        self.convs = nn.ModuleList()
        self.emlps = nn.ModuleList()
        self.batch_norms = nn.ModuleList()
        for _ in range(self.num_gnn_layers):
            kw = dict(in_channels=n_hidden, out_channels=n_hidden, aggregators=aggregators, scalers=scalers, deg=deg,
                      edge_dim=n_hidden, towers=1, pre_layers=1, post_layers=1, divide_input=False)
            conv = PNAConvHetero(n_hidden=n_hidden, **kw) if reverse_mp else PNAConv(**kw)
            if self.edge_updates:
                self.emlps.append(nn.Sequential(nn.Linear(3 * n_hidden, n_hidden), nn.ReLU(),
                                                nn.Linear(n_hidden, n_hidden)))
            self.convs.append(conv)
            self.batch_norms.append(BatchNorm(n_hidden))

    def forward(self, x, edge_index, edge_attr):
        x = _embed(self.node_emb, x.view(x.shape[0], -1))
        e = _embed(self.edge_emb, edge_attr.view(edge_attr.shape[0], -1))
        return self._layers(x, edge_index, e)


# ----------------------------------------------------------------------------- per-column PNA
class _ColumnStacks(nn.Module):
    """Shared body of ``CPNA`` (/root/reference/src/nn/gnn/pna.py:164-231) and ``CPNATAB`` (:233-304): one PNA stack per
    edge COLUMN -- ``col_convs[c][i]``, ``col_batch_norms[c][i]``, ``col_emlps[c][i]`` -- run column after column over the
    SAME node state x (column c + 1 starts from the x column c produced, pna.py:219-230), each updating only its own
    [E, H] column of edge_attr.  The CSC structure is built once for all K * L conv calls; every column is read from and
    written back to the caller's [E, K, H] layout through the kernels' row maps (no permutation passes for H = 64)."""

    def _init_stacks(self, num_features, num_gnn_layers, n_hidden, edge_updates, edge_dim, final_dropout, deg, reverse_mp):
        self.n_hidden = n_hidden
        self.num_gnn_layers = num_gnn_layers
        self.edge_updates = edge_updates
        self.final_dropout = final_dropout
        self.reverse_mp = reverse_mp
        self.num_cols = edge_dim // n_hidden
        aggregators = ['mean', 'min', 'max', 'std']
        scalers = ['identity', 'amplification', 'attenuation']
        self.node_emb = nn.Linear(num_features, n_hidden)
        self.col_convs = nn.ModuleList()
        self.col_emlps = nn.ModuleList()
        self.col_batch_norms = nn.ModuleList()
        for _ in range(self.num_cols):
            convs, emlps, batch_norms = nn.ModuleList(), nn.ModuleList(), nn.ModuleList()
            for _ in range(num_gnn_layers):
                kw = dict(in_channels=n_hidden, out_channels=n_hidden, aggregators=aggregators, scalers=scalers, deg=deg,
                          edge_dim=n_hidden, towers=1, pre_layers=1, post_layers=1, divide_input=False)
                convs.append(PNAConvHetero(n_hidden=n_hidden, **kw) if reverse_mp else PNAConv(**kw))
                if edge_updates:
                    emlps.append(nn.Sequential(nn.Linear(3 * n_hidden, n_hidden), nn.ReLU(), nn.Linear(n_hidden, n_hidden)))
                batch_norms.append(BatchNorm(n_hidden))
            self.col_convs.append(convs)
            self.col_emlps.append(emlps)
            self.col_batch_norms.append(batch_norms)

    def _columns(self, x, edge_index, edge_attr):
        """-> (x, [K] list of updated [E, H] columns in the caller's edge order)"""
        x = self.node_emb(x.view(x.shape[0], -1))
        st = edge_structure(edge_index, x.shape[0], reverse=self.reverse_mp)
        L = self.num_gnn_layers
        fast = (not self.reverse_mp) and (x.shape[1] == 64 or BF.wide_tc_available(x.shape[1])) and BF.row_maps_available()
        perm32 = st.fwd.perm32() if fast else None
        cols = []
        for c in range(self.num_cols):
            e_in = e = edge_attr[:, c, :].contiguous()
            if not fast:
                e = BF.gather_rows(e, st.fwd.perm)
            for i in range(L):
                sink = BF.EdgeGradSink() if (self.edge_updates and e.requires_grad and torch.is_grad_enabled()) else None
                bn = self.col_batch_norms[c][i].module
                if fast:
                    e_eid = perm32 if (i == 0 or not self.edge_updates) else None
                    out_eid = perm32 if i == L - 1 else None
                    x = BF.bn_relu_residual(bn, self.col_convs[c][i].forward_sorted(x, st.fwd, e, sink=sink, e_eid=e_eid), x)
                    if self.edge_updates:
                        e = edge_update_sorted(self.col_emlps[c][i], x, st, e, 1.0, 0.5, sink=sink, e_eid=e_eid, out_eid=out_eid)
                else:
                    x = BF.bn_relu_residual(bn, _conv_sorted(self.col_convs[c][i], x, st, e, sink), x)
                    if self.edge_updates:
                        e = edge_update_sorted(self.col_emlps[c][i], x, st, e, 1.0, 0.5, sink=sink)
            if not self.edge_updates:
                e = e_in
            elif not fast:
                e = BF.scatter_rows(e, st.fwd.perm)
            cols.append(e)
        return x, cols


class CPNA(_ColumnStacks):
    def __init__(self, num_features, num_gnn_layers, n_classes=2, n_hidden=128, edge_updates=True, edge_dim=None, dropout=0.0,
                 final_dropout=0.5, deg=None, reverse_mp=False):
        super().__init__()
        self._init_stacks(num_features, num_gnn_layers, n_hidden, edge_updates, edge_dim, final_dropout, deg, reverse_mp)

    def forward(self, x, edge_index, edge_attr):
        x, cols = self._columns(x, edge_index, edge_attr)
        return x, torch.stack(cols, dim=1)          # the reference writes the columns back in place (pna.py:230)


class CPNATAB(_ColumnStacks):
    def __init__(self, num_features, num_gnn_layers, n_classes=2, n_hidden=128, edge_updates=True, edge_dim=None, dropout=0.0,
                 final_dropout=0.5, deg=None, reverse_mp=False):
        super().__init__()
        self._init_stacks(num_features, num_gnn_layers, n_hidden, edge_updates, edge_dim, final_dropout, deg, reverse_mp)
        self.row_atts = nn.ModuleList()
        self.row_norms = nn.ModuleList()
        for _ in range(self.num_cols):               # one per COLUMN, indexed by LAYER in forward: pna.py:279-280, 299-300
            self.row_norms.append(nn.LayerNorm(n_hidden))
            self.row_atts.append(nn.TransformerEncoderLayer(n_hidden, nhead=8, batch_first=True))

    def forward(self, x, edge_index, edge_attr):
        x, cols = self._columns(x, edge_index, edge_attr)
        tok = torch.stack(cols, dim=1)
        for i in range(self.num_gnn_layers):         # attention over the K column tokens of every edge row (pna.py:299-300)
            att = BF.transformer_encoder_layer(self.row_atts[i], tok)
            tok = (tok + BF.add_dropout_layernorm(att, None, 0.0, self.row_norms[i])) / 2
        return x, tok
