"""Drop-ins for the tabular+graph backbones: ``TABGNN`` (/root/reference/src/nn/models/tabgnn.py:27-219),
``TABGNNInterleaved`` (src/nn/models/inteleaved.py:28-228) and ``TABGNNFused``
(src/nn/models/fused.py:31-269).  Same constructors, forward signatures, state_dict keys
 and residual conventions (note the differing parenthesisation of the column
attention residual: tabgnn.py:219 vs fused.py:249 / inteleaved.py:219).

Graph side (PNAConv -> BatchNorm -> ReLU -> residual, edge update) runs on the b200mp CUDA kernels
with the edge tensors held in CSC storage order for all layers.  The column attention over the
S = K + 1 tokens of a row keeps ``nn.TransformerEncoderLayer`` as the parameter container (same attribute
names ``tab_conv`` / ``tab_norm``, same state_dict keys) and runs it through
``functional.transformer_encoder_layer``: projections on the tcgen05 node kernels, attention core on
csrc/col_attention.cu; it is row-wise, hence order-independent.
"""
from __future__ import annotations

from typing import Optional

import os as _os

import torch
import torch.nn.functional as F
from torch import Tensor, nn

from .. import column_stack as CS
from .. import functional as BF
from ..graph import EdgeStructure, edge_structure
from .backbones import PNAConvHetero, _EdgeStorage, _embed_wide
from .conv import BatchNorm, PNAConv, edge_update_sorted

_AGGR = ['mean', 'max', 'min', 'std']
_SCAL = ['identity', 'amplification', 'attenuation']


def _xavier(module: nn.Module):
    for p in module.parameters():
        if p.dim() > 1:
            nn.init.xavier_uniform_(p)


class _ColumnAttention(nn.Module):
    """``tab_norm(tab_conv(t))`` over [rows, S, C] tokens."""

    def _init_attention(self, channels, nhead, feedforward_channels, dropout, activation):
        self.tab_conv = nn.TransformerEncoderLayer(d_model=channels, nhead=nhead,
                                                   dim_feedforward=feedforward_channels or channels, dropout=dropout,
                                                   activation=activation, batch_first=True)
        self.tab_norm = nn.LayerNorm(channels)

    #: rows per chunk of the recomputing evaluation (functional.row_chunked): above this many token rows the layer is run
    #: chunk by chunk and re-run in backward, so that only one chunk's ~12 intermediate [rows, S, C] tensors are alive
    chunk_rows = 1 << 20

    def _attend_rows(self, t: Tensor) -> Tensor:
        # tab_conv holds the parameters (state_dict keys of nn.TransformerEncoderLayer); the math runs on
        # the b200mp kernels (functional.transformer_encoder_layer)
        return BF.add_dropout_layernorm(BF.transformer_encoder_layer(self.tab_conv, t), None, 0.0, self.tab_norm)

    def attend(self, t: Tensor) -> Tensor:
        params = [*self.tab_conv.parameters(), *self.tab_norm.parameters()]
        return BF.row_chunked(self._attend_rows, t, params, self.chunk_rows if torch.is_grad_enabled() else 0)

    def _reset_attention(self):
        _xavier(self.tab_conv)
        self.tab_norm.reset_parameters()


class _GraphStep(nn.Module):
    """gnn_conv / gnn_norm / gnn_edge_update of one layer, on sorted (CSC-order) edge tensors."""

    def _init_graph(self, nhidden, deg, reverse_mp):
        kw = dict(in_channels=nhidden, out_channels=nhidden, aggregators=_AGGR, scalers=_SCAL, deg=deg, edge_dim=nhidden,
                  towers=1, pre_layers=1, post_layers=1, divide_input=False)
        self.gnn_conv = PNAConvHetero(n_hidden=nhidden, **kw) if reverse_mp else PNAConv(**kw)
        self.gnn_norm = BatchNorm(nhidden)
        self.gnn_edge_update = nn.Sequential(nn.Linear(3 * nhidden, nhidden), nn.ReLU(), nn.Linear(nhidden, nhidden))

    def _reset_graph(self):
        self.gnn_conv.reset_parameters()
        self.gnn_norm.reset_parameters()
        _xavier(self.gnn_edge_update)

    def graph_step(self, x: Tensor, st: EdgeStructure, e: Tensor, alpha: float, beta: float, part=None, e_eid=None, out_eid=None):
        """part (parallel.Partition): x = the owned node rows, e = the owned edges in local CSC order; source-side node tables
        are all-gathered, BatchNorm statistics synchronised (as _Backbone.message_passing_partitioned).  e_eid / out_eid: int32
        row maps when e comes in / e' goes out in the caller's edge order instead of CSC order (H = 64 kernels).  Conv and
        edge update read the same e: their gradients meet in one buffer (functional.EdgeGradSink)."""
        sink = BF.EdgeGradSink() if (e.requires_grad and torch.is_grad_enabled()) else None
        if part is None:
            c = self.gnn_conv.forward_sorted(x, st, e, sink=sink) if isinstance(self.gnn_conv, PNAConvHetero) \
                else self.gnn_conv.forward_sorted(x, st.fwd, e, sink=sink, e_eid=e_eid)
            x = BF.bn_relu_residual(self.gnn_norm.module, c, x)
            return x, edge_update_sorted(self.gnn_edge_update, x, st, e, alpha, beta, sink=sink, e_eid=e_eid, out_eid=out_eid)
        from .. import parallel as PL
        if isinstance(self.gnn_conv, PNAConvHetero):
            raise NotImplementedError("reverse_mp is not supported on a partitioned graph")
        gather = lambda t: PL.all_gather_rows(t, part)
        c = self.gnn_conv.forward_sorted(x, part.st.fwd, e, gather=gather, sink=sink)
        bn = self.gnn_norm.module
        if bn.training:
            x = BF.bn_relu_residual(bn, c, x, group=part.group, n_total=part.num_nodes)
        else:
            x = (x + F.relu(PL.sync_batch_norm(bn, c, part))) / 2
        return x, edge_update_sorted(self.gnn_edge_update, x, part.st, e, alpha, beta, gather=gather, sink=sink)


class _SharedParams:
    def zero_grad_shared_params(self):
        for p in self.get_shared_params():
            if p.grad is not None:
                p.grad.data.zero_()


# ----------------------------------------------------------------------------- TABGNN
class FTTransformerLayer(_ColumnAttention):
    def __init__(self, channels, nhead, feedforward_channels=None, dropout=0.5, activation='relu', nhidden=128):
        super().__init__()
        self.channels, self.nhidden = channels, nhidden
        self._init_attention(channels, nhead, feedforward_channels, dropout, activation)
        self.reset_parameters()

    def reset_parameters(self):
        self._reset_attention()

    def forward(self, x_tab):
        return (x_tab + self.attend(x_tab)) / 2


class PNALayer(_GraphStep):
    def __init__(self, channels, nhidden=128, deg=None, reverse_mp=False):
        super().__init__()
        self.channels, self.nhidden, self.reverse_mp = channels, nhidden, reverse_mp
        self._init_graph(nhidden, deg, reverse_mp)
        self.reset_parameters()

    def reset_parameters(self):
        self._reset_graph()

    def forward_sorted(self, x, st, e, part=None, e_eid=None, out_eid=None):
        return self.graph_step(x, st, e, 1.0, 0.5, part, e_eid, out_eid)            # e + emlp(.)/2  (tabgnn.py:190)

    def forward(self, x_gnn, edge_index, edge_attr):
        st = edge_structure(edge_index, x_gnn.shape[0], reverse=self.reverse_mp)
        x, e = self.forward_sorted(x_gnn, st, BF.gather_rows(edge_attr, st.fwd.perm))
        return x, BF.scatter_rows(e, st.fwd.perm)


class TABGNN(_EdgeStorage, _SharedParams, nn.Module):
    def __init__(self, channels: int, num_layers: int, deg=None, node_dim: int = 1, nhidden: int = 128,
                 edge_dim: int = None, reverse_mp: bool = False, feedforward_channels: Optional[int] = None,
                 nhead: int = 8, dropout: float = 0.5, activation: str = 'relu') -> None:
        super().__init__()
        if num_layers <= 0:
            raise ValueError(f"num_layers must be a positive integer (got {num_layers})")
        self.channels, self.nhidden, self.reverse_mp = channels, nhidden, reverse_mp
        self.node_dim, self.edge_dim = node_dim + channels, edge_dim + channels
        self.cls_embedding = nn.Parameter(torch.empty(channels))
        self.node_emb = nn.Linear(self.node_dim, nhidden)
        self.edge_emb = nn.Linear(self.edge_dim, nhidden)
        self.tabular_backbone = nn.ModuleList(
            FTTransformerLayer(channels, nhead, feedforward_channels, dropout, activation, nhidden) for _ in range(num_layers))
        self.gnn_backbone = nn.ModuleList(PNALayer(channels, nhidden, deg, reverse_mp) for _ in range(num_layers))
        self.reset_parameters()

    def reset_parameters(self):
        nn.init.normal_(self.cls_embedding, std=0.01)
        self.node_emb.reset_parameters()
        self.edge_emb.reset_parameters()
        for layer in list(self.tabular_backbone) + list(self.gnn_backbone):
            layer.reset_parameters()

    def get_shared_params(self):
        return [self.cls_embedding, *self.node_emb.parameters(), *self.edge_emb.parameters(),
                *self.tabular_backbone.parameters(), *self.gnn_backbone.parameters()]

    def _with_cls(self, t: Tensor) -> Tensor:
        return torch.cat([self.cls_embedding.expand(t.shape[0], 1, -1), t], dim=1)

    def _tabular(self, t: Tensor) -> Tensor:
        """t + T_L(..T_1(t)), T_l(u) = (u + tab_norm_l(tab_conv_l(u))) / 2 (tabgnn.py:116-131; the same layers serve node and edge
        tokens): one autograd node over the whole stack where the kernels cover the shapes (column_stack.supported), else layer by layer"""
        convs = [layer.tab_conv for layer in self.tabular_backbone]
        norms = [layer.tab_norm for layer in self.tabular_backbone]
        if CS.supported(convs, norms, t) and _os.environ.get("B200MP_TAB_STACK", "1") != "0":
            return CS.column_stack(convs, norms, t, self.tabular_backbone[0].chunk_rows, add_input=True)
        u = t
        for layer in self.tabular_backbone:
            u = layer(u)
        return t + u

    def forward(self, x, edge_index, edge_attr):
        x, e = self._with_cls(x), self._with_cls(edge_attr)
        sx, se = self._tabular(x), self._tabular(e)           # x + x_tab, e + e_tab
        x = _embed_wide(self.node_emb, sx.reshape(x.shape[0], -1), 0.5)
        e = self._store(_embed_wide(self.edge_emb, se.reshape(e.shape[0], -1), 0.5))
        st = edge_structure(edge_index, x.shape[0], reverse=self.reverse_mp)
        if not self.reverse_mp and x.shape[1] == 64 and BF.row_maps_available():
            # e stays in the caller's order at both ends: first layer reads / last layer writes through the CSC permutation
            perm32, L = st.fwd.perm32(), len(self.gnn_backbone)
            for i, layer in enumerate(self.gnn_backbone):
                x, e = layer.forward_sorted(x, st, e, e_eid=perm32 if i == 0 else None, out_eid=perm32 if i == L - 1 else None)
            return x, e
        e = BF.gather_rows(e, st.fwd.perm)
        for layer in self.gnn_backbone:
            x, e = layer.forward_sorted(x, st, e)
        return x, BF.scatter_rows(e, st.fwd.perm)

    def forward_partitioned(self, x_local, edge_attr_local, part):
        """One rank of a destination-range partition (parallel.Partition): x_local = token rows of the owned accounts,
        edge_attr_local = token rows of the owned transactions in local CSC order (part.shard_edges).  The column attention is
        row-local, so it shards with the rows; the graph half exchanges source-side node tables per layer.
        Returns the owned rows of (x, edge_attr) in the same orders."""
        x, e = self._with_cls(x_local), self._with_cls(edge_attr_local)
        sx, se = self._tabular(x), self._tabular(e)
        x = _embed_wide(self.node_emb, sx.reshape(x.shape[0], -1), 0.5)
        e = self._store(_embed_wide(self.edge_emb, se.reshape(e.shape[0], -1), 0.5))
        for layer in self.gnn_backbone:
            x, e = layer.forward_sorted(x, None, e, part)
        return x, e


# ----------------------------------------------------------------------------- shared top level of the two mixed models
class _TabGraphModel(_SharedParams, _ColumnAttention):
    """cls_embedding, node_emb, edge_emb, top-level tab_conv / tab_norm, backbone[...]."""

    def _init_top(self, layer_cls, channels, num_layers, encoder, deg, node_dim, nhidden, edge_dim, reverse_mp,
                  feedforward_channels, nhead, dropout, activation):
        if num_layers <= 0:
            raise ValueError(f"num_layers must be a positive integer (got {num_layers})")
        self.channels, self.nhidden, self.node_dim = channels, nhidden, node_dim
        self.edge_dim, self.encoder, self.reverse_mp = edge_dim + channels, encoder, reverse_mp
        self.cls_embedding = nn.Parameter(torch.empty(channels))
        self.node_emb = nn.Linear(node_dim, nhidden)
        self.edge_emb = nn.Linear(self.edge_dim, nhidden)
        self._init_attention(channels, nhead, feedforward_channels, dropout, activation)
        self.backbone = nn.ModuleList(
            layer_cls(channels, nhead, feedforward_channels, dropout, activation, nhidden, deg, reverse_mp)
            for _ in range(num_layers))
        self.reset_parameters()

    def reset_parameters(self):
        nn.init.normal_(self.cls_embedding, std=0.01)
        self.node_emb.reset_parameters()
        self.edge_emb.reset_parameters()
        self._reset_attention()
        for layer in self.backbone:
            layer.reset_parameters()

    def get_shared_params(self):
        enc = list(self.encoder.parameters()) if isinstance(self.encoder, nn.Module) else []
        return [*enc, *self.tab_conv.parameters(), *self.tab_norm.parameters(), self.cls_embedding,
                *self.node_emb.parameters(), *self.edge_emb.parameters(), *self.backbone.parameters()]

    def _with_cls(self, t: Tensor) -> Tensor:
        return torch.cat([self.cls_embedding.expand(t.shape[0], 1, -1), t], dim=1)


class _MixedLayer(_ColumnAttention, _GraphStep):
    def _init_mixed(self, channels, nhead, feedforward_channels, dropout, activation, nhidden, deg, reverse_mp):
        self.channels, self.nhidden = channels, nhidden
        self._init_attention(channels, nhead, feedforward_channels, dropout, activation)
        self._init_graph(nhidden, deg, reverse_mp)


# ----------------------------------------------------------------------------- TABGNNInterleaved
class FTTransformerPNAInterleavedLayer(_MixedLayer):
    def __init__(self, channels, nhead, feedforward_channels=None, dropout=0.5, activation='relu', nhidden=128,
                 deg=None, reverse_mp=False):
        super().__init__()
        self._init_mixed(channels, nhead, feedforward_channels, dropout, activation, nhidden, deg, reverse_mp)
        self.reset_parameters()

    def reset_parameters(self):
        self._reset_attention()
        self._reset_graph()

    def forward_sorted(self, x, st, tok, part=None):
        """tok [E,S,C] in CSC order; the CLS token is the edge feature of the PNA step."""
        tok = tok + self.attend(tok) / 2                              # inteleaved.py:219
        x, cls = self.graph_step(x, st, tok[:, 0, :].contiguous(), 0.5, 0.5, part)   # (e + emlp(.))/2, :225
        return x, torch.cat([cls.unsqueeze(1), tok[:, 1:, :]], dim=1)

    def forward(self, x_gnn, edge_index, edge_attr):
        st = edge_structure(edge_index, x_gnn.shape[0], reverse=isinstance(self.gnn_conv, PNAConvHetero))
        E, S, C = edge_attr.shape
        tok = BF.gather_rows(edge_attr.reshape(E, S * C), st.fwd.perm).view(E, S, C)
        x, tok = self.forward_sorted(x_gnn, st, tok)
        return x, BF.scatter_rows(tok.reshape(E, S * C), st.fwd.perm).view(E, S, C)


class TABGNNInterleaved(_TabGraphModel):
    def __init__(self, channels: int, num_layers: int, encoder=None, deg=None, node_dim: int = 1, nhidden: int = 128,
                 edge_dim: int = None, reverse_mp: bool = False, feedforward_channels: Optional[int] = None,
                 nhead: int = 8, dropout: float = 0.5, activation: str = 'relu') -> None:
        super().__init__()
        if channels != nhidden:
            raise ValueError("TABGNNInterleaved feeds the [E, channels] CLS token to a PNAConv with edge_dim=nhidden: "
                             "channels must equal nhidden")
        self._init_top(FTTransformerPNAInterleavedLayer, channels, num_layers, encoder, deg, node_dim, nhidden, edge_dim,
                       reverse_mp, feedforward_channels, nhead, dropout, activation)

    def forward(self, x, edge_index, edge_attr):
        x = self.node_emb(x.reshape(-1, self.node_dim))
        tok0 = self._with_cls(edge_attr)
        tok0 = (tok0 + self.attend(tok0)) / 2                         # inteleaved.py:155
        st = edge_structure(edge_index, x.shape[0], reverse=self.reverse_mp)
        E, S, C = tok0.shape
        tok0 = BF.gather_rows(tok0.reshape(E, S * C), st.fwd.perm).view(E, S, C)
        tok = tok0
        for layer in self.backbone:
            x, tok = layer.forward_sorted(x, st, tok)
        cls = ((tok + tok0) / 2)[:, 0, :].contiguous()               # :161-162
        return x, BF.scatter_rows(cls, st.fwd.perm)

    def forward_partitioned(self, x_local, edge_attr_local, part):
        """as TABGNN.forward_partitioned: owned node rows / owned edge token rows in local CSC order in, the same out"""
        x = self.node_emb(x_local.reshape(-1, self.node_dim))
        tok0 = self._with_cls(edge_attr_local)
        tok0 = (tok0 + self.attend(tok0)) / 2
        tok = tok0
        for layer in self.backbone:
            x, tok = layer.forward_sorted(x, None, tok, part)
        return x, ((tok + tok0) / 2)[:, 0, :].contiguous()


# ----------------------------------------------------------------------------- TABGNNFused
class FTTransformerPNAFusedLayer(_MixedLayer):
    def __init__(self, channels, nhead, feedforward_channels=None, dropout=0.5, activation='relu', nhidden=128,
                 deg=None, reverse_mp=False):
        super().__init__()
        self._init_mixed(channels, nhead, feedforward_channels, dropout, activation, nhidden, deg, reverse_mp)
        D = channels + 2 * nhidden
        self.fuse = nn.Sequential(nn.LayerNorm(D), nn.Linear(D, 4 * D), nn.LeakyReLU(), nn.Dropout(dropout),
                                  nn.Linear(4 * D, 4 * D), nn.LeakyReLU(), nn.Dropout(dropout), nn.Linear(4 * D, D))
        self.fuse_norm = nn.LayerNorm(D)
        self.reset_parameters()

    def reset_parameters(self):
        self._reset_attention()
        self._reset_graph()
        _xavier(self.fuse)
        self.fuse_norm.reset_parameters()

    def forward_sorted(self, x_tab, x, st, e, target_edge_index, lp):
        C, H = self.channels, self.nhidden
        x_tab = x_tab + self.attend(x_tab) / 2                        # fused.py:249
        x, e = self.graph_step(x, st, e, 0.5, 0.5)                    # :252-254
        if not lp:                                                    # :256-268, restated without unique / in-place
            cls = x_tab[:, 0, :]
            z = torch.cat([cls, x[target_edge_index[0]], x[target_edge_index[1]]], dim=-1)
            z = (z + self.fuse_norm(self.fuse(z))) / 2
            x_tab = torch.cat([((cls + z[:, :C]) / 2).unsqueeze(1), x_tab[:, 1:, :]], dim=1)
            # mean of the target-edge embeddings per node, averaged into x (:260-268): one kernel pass, no unique / atomics
            x = BF.segment_mean_blend(x, torch.cat([z[:, C:C + H], z[:, C + H:]], dim=0), target_edge_index.reshape(-1))
        return x_tab, x, e

    def forward(self, x_tab, x_gnn, edge_index, edge_attr, target_edge_index, lp=False):
        st = edge_structure(edge_index, x_gnn.shape[0], reverse=isinstance(self.gnn_conv, PNAConvHetero))
        x_tab, x, e = self.forward_sorted(x_tab, x_gnn, st, BF.gather_rows(edge_attr, st.fwd.perm), target_edge_index, lp)
        return x_tab, x, BF.scatter_rows(e, st.fwd.perm)


class TABGNNFused(_TabGraphModel):
    def __init__(self, channels: int, num_layers: int, encoder=None, deg=None, node_dim: int = 1, nhidden: int = 128,
                 edge_dim: int = None, reverse_mp: bool = False, feedforward_channels: Optional[int] = None,
                 nhead: int = 8, dropout: float = 0.5, activation: str = 'relu') -> None:
        super().__init__()
        self._init_top(FTTransformerPNAFusedLayer, channels, num_layers, encoder, deg, node_dim, nhidden, edge_dim,
                       reverse_mp, feedforward_channels, nhead, dropout, activation)

    def forward(self, x, edge_index, edge_attr, target_edge_index, target_edge_attr, lp=False):
        x = self.node_emb(x.reshape(-1, self.node_dim))
        tgt0 = self.attend(self._with_cls(target_edge_attr))          # fused.py:160 (no residual)
        tok = self._with_cls(edge_attr)
        tok = (tok + self.attend(tok)) / 2                            # :164
        e = _embed_wide(self.edge_emb, tok.reshape(tok.shape[0], -1))
        st = edge_structure(edge_index, x.shape[0], reverse=self.reverse_mp)
        e = BF.gather_rows(e, st.fwd.perm)
        x_tab = tgt0
        for layer in self.backbone:
            x_tab, x, e = layer.forward_sorted(x_tab, x, st, e, target_edge_index, lp)
        tgt = _embed_wide(self.edge_emb, (x_tab + tgt0).reshape(tgt0.shape[0], -1), 0.5)   # :172-174
        return x, BF.scatter_rows(e, st.fwd.perm), tgt
