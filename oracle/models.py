"""ORACLE (test infrastructure, not product code): functional CPU restatement of the
reference's backbones over a plain state_dict.  Each function cites the reference lines it
follows; arithmetic of the PyG operators comes from oracle/pyg_functional.py.

Pinned by tests/test_oracle_golden.py against fixtures in tests/golden/ that were produced by
the reference's own, unmodified model files (oracle/gen_golden.py).  The PyG operators under
those files are themselves restated => "parity unpinned" at that boundary.
"""
from __future__ import annotations

from typing import Dict, Sequence

import torch
import torch.nn.functional as F
from torch import Tensor

from . import pyg_functional as PF
from .pyg_functional import sub

Params = Dict[str, Tensor]

AGG_PNAS = ("mean", "min", "max", "std")       # src/nn/gnn/pna.py:60
AGG_TAB = ("mean", "max", "min", "std")        # src/nn/models/{tabgnn.py:160,fused.py:200,inteleaved.py:185}
SCALERS = ("identity", "amplification", "attenuation")


def cast(p: Params, dtype) -> Params:
    return {k: (v.to(dtype) if v.is_floating_point() else v.clone()) for k, v in p.items()}


def _lin(p: Params, name: str, x: Tensor) -> Tensor:
    return F.linear(x, p[name + ".weight"], p[name + ".bias"])


def _emlp(p: Params, prefix: str, x: Tensor) -> Tensor:
    """Sequential(Linear(3H,H), ReLU, Linear(H,H)) - gine.py:71-75, pna.py:80-84."""
    return _lin(p, prefix + "2", _lin(p, prefix + "0", x).relu())


def _conv(p: Params, prefix: str, kind: str, reverse_mp: bool, x, edge_index, edge_attr, aggregators):
    q = sub(p, prefix)
    if kind == "gine":
        if reverse_mp:   # gine.py:28-35: (x, None) => no self term
            return PF.hetero(PF.gine_conv, q, x, edge_index, edge_attr, self_term=False)
        return PF.gine_conv(q, x, edge_index, edge_attr)
    if reverse_mp:       # pna.py:37-46
        return PF.hetero(PF.pna_conv, q, x, edge_index, edge_attr, aggregators=aggregators, scalers=SCALERS)
    return PF.pna_conv(q, x, edge_index, edge_attr, aggregators, SCALERS)


def mp_stack(p: Params, kind: str, x: Tensor, edge_index: Tensor, e: Tensor, num_layers: int,
             edge_updates: bool, reverse_mp: bool = False, training: bool = True,
             aggregators: Sequence[str] = AGG_PNAS, update_mode: str = "add_half"):
    """The L-layer message-passing loop of gine.py:86-89 / pna.py:95-99 on already-embedded
    x[N,H], e[E,H] (this is the region BASELINE.json's metric times):
        x = (x + relu(BN(conv(x, ei, e)))) / 2 ;  e = e + emlp([x[src], x[dst], e]) / 2."""
    src, dst = edge_index[0], edge_index[1]
    for i in range(num_layers):
        h = _conv(p, f"convs.{i}.", kind, reverse_mp, x, edge_index, e, aggregators)
        h = PF.batch_norm(sub_inplace(p, f"batch_norms.{i}."), h, training)
        x = (x + F.relu(h)) / 2
        if edge_updates:
            e = e + _emlp(p, f"emlps.{i}.", torch.cat([x[src], x[dst], e], dim=-1)) / 2
    return x, e


class sub_inplace(dict):
    """Prefix view that writes through to the parent dict (BatchNorm running stats)."""

    def __init__(self, parent: Params, prefix: str):
        super().__init__()
        self._parent, self._prefix = parent, prefix

    def __getitem__(self, k):
        return self._parent[self._prefix + k]

    def __contains__(self, k):
        return (self._prefix + k) in self._parent

    def get(self, k, default=None):
        return self._parent.get(self._prefix + k, default)

    def __setitem__(self, k, v):
        self._parent[self._prefix + k] = v


def gine(p: Params, x: Tensor, edge_index: Tensor, edge_attr: Tensor, num_layers: int,
         edge_updates: bool, reverse_mp: bool = False, training: bool = True):
    """GINe.forward, src/nn/gnn/gine.py:80-91."""
    x = _lin(p, "node_emb", x)                                       # :83
    e = _lin(p, "edge_emb", edge_attr.reshape(edge_attr.shape[0], -1))   # :84
    return mp_stack(p, "gine", x, edge_index, e, num_layers, edge_updates, reverse_mp, training)


def pnas(p: Params, x: Tensor, edge_index: Tensor, edge_attr: Tensor, num_layers: int,
         edge_updates: bool, reverse_mp: bool = False, training: bool = True):
    """PNAS.forward, src/nn/gnn/pna.py:90-101."""
    x = _lin(p, "node_emb", x.reshape(x.shape[0], -1))               # :91
    e = _lin(p, "edge_emb", edge_attr.reshape(edge_attr.shape[0], -1))   # :93
    return mp_stack(p, "pna", x, edge_index, e, num_layers, edge_updates, reverse_mp, training,
                    aggregators=AGG_PNAS)


# ----------------------------------------------------------------- column attention (a12)
def transformer_encoder_layer(p: Params, prefix: str, x: Tensor, nhead: int) -> Tensor:
    """torch.nn.TransformerEncoderLayer(d, nhead, dim_ff, dropout=0 / eval, relu, post-norm,
    batch_first) restated with explicit math-SDP (main.py:27-29 forces the math kernel).
    Keys: self_attn.in_proj_{weight,bias}, self_attn.out_proj.*, linear1.*, linear2.*, norm1.*, norm2.*"""
    q = sub(p, prefix)
    R, S, C = x.shape
    hd = C // nhead
    qkv = F.linear(x, q["self_attn.in_proj_weight"], q["self_attn.in_proj_bias"])
    qh, kh, vh = [t.reshape(R, S, nhead, hd).transpose(1, 2) for t in qkv.chunk(3, dim=-1)]
    att = torch.softmax((qh @ kh.transpose(-1, -2)) / (hd ** 0.5), dim=-1)
    o = (att @ vh).transpose(1, 2).reshape(R, S, C)
    o = F.linear(o, q["self_attn.out_proj.weight"], q["self_attn.out_proj.bias"])
    x = F.layer_norm(x + o, (C,), q["norm1.weight"], q["norm1.bias"])
    f = F.linear(F.linear(x, q["linear1.weight"], q["linear1.bias"]).relu(),
                 q["linear2.weight"], q["linear2.bias"])
    return F.layer_norm(x + f, (C,), q["norm2.weight"], q["norm2.bias"])


def tab_layer(p: Params, prefix: str, x: Tensor, nhead: int) -> Tensor:
    """tab_norm(tab_conv(x)): the unit every variant wraps with its own residual."""
    t = transformer_encoder_layer(p, prefix + "tab_conv.", x, nhead)
    C = x.shape[-1]
    return F.layer_norm(t, (C,), p[prefix + "tab_norm.weight"], p[prefix + "tab_norm.bias"])


def _pna_layer(p: Params, prefix: str, x, edge_index, e, reverse_mp, training, mode):
    """PNALayer.forward (models/tabgnn.py:187-191) / the PNA half of the fused & interleaved
    layers (models/fused.py:252-254, inteleaved.py:222-225).  mode 'add_half': e + mlp/2 ;
    'avg': (e + mlp)/2."""
    src, dst = edge_index[0], edge_index[1]
    h = _conv(p, prefix + "gnn_conv.", "pna", reverse_mp, x, edge_index, e, AGG_TAB)
    h = PF.batch_norm(sub_inplace(p, prefix + "gnn_norm."), h, training)
    x = (x + F.relu(h)) / 2
    upd = _emlp(p, prefix + "gnn_edge_update.", torch.cat([x[src], x[dst], e], dim=-1))
    e = e + upd / 2 if mode == "add_half" else (e + upd) / 2
    return x, e


def tabgnn(p: Params, x: Tensor, edge_index: Tensor, edge_attr: Tensor, num_layers: int,
           nhead: int = 8, reverse_mp: bool = False, training: bool = True):
    """TABGNN.forward, src/nn/models/tabgnn.py:101-151 (dropout must be 0 / eval)."""
    V, Ne = x.shape[0], edge_attr.shape[0]
    cls = p["cls_embedding"]
    x = torch.cat([cls.expand(V, 1, -1), x], dim=1)                  # :115-117
    edge_attr = torch.cat([cls.expand(Ne, 1, -1), edge_attr], dim=1)     # :120-121
    t_x, t_e = x, edge_attr
    for i in range(num_layers):                                      # :126-129
        pre = f"tabular_backbone.{i}."
        t_x = (t_x + tab_layer(p, pre, t_x, nhead)) / 2              # FTTransformerLayer :218-219
        t_e = (t_e + tab_layer(p, pre, t_e, nhead)) / 2
    x = (x + t_x) / 2                                                # :132-133
    edge_attr = (edge_attr + t_e) / 2
    x = _lin(p, "node_emb", x.reshape(V, -1))                        # :136-137
    e = _lin(p, "edge_emb", edge_attr.reshape(Ne, -1))               # :140-141
    for i in range(num_layers):                                      # :147-148
        x, e = _pna_layer(p, f"gnn_backbone.{i}.", x, edge_index, e, reverse_mp, training, "add_half")
    return x, e


def tabgnn_interleaved(p: Params, x: Tensor, edge_index: Tensor, edge_attr: Tensor, num_layers: int,
                       node_dim: int, nhead: int = 8, reverse_mp: bool = False, training: bool = True):
    """TABGNNInterleaved.forward, src/nn/models/inteleaved.py:141-163, layer :218-228."""
    Ne = edge_attr.shape[0]
    x = _lin(p, "node_emb", x.reshape(-1, node_dim))
    ea = torch.cat([p["cls_embedding"].expand(Ne, 1, -1), edge_attr], dim=1)
    ea = (ea + tab_layer(p, "", ea, nhead)) / 2                      # :155
    e_attr = ea
    for i in range(num_layers):
        pre = f"backbone.{i}."
        e_attr = e_attr + tab_layer(p, pre, e_attr, nhead) / 2       # :219  (x + LN(T(x))/2)
        cls_tok, feat = e_attr[:, 0, :], e_attr[:, 1:, :]
        x, cls_tok = _pna_layer(p, pre, x, edge_index, cls_tok, reverse_mp, training, "avg")
        e_attr = torch.cat([cls_tok.unsqueeze(1), feat], dim=1)
    ea = (e_attr + ea) / 2                                           # :161
    return x, ea[:, 0, :]


def _fuse_mlp(p: Params, prefix: str, x: Tensor) -> Tensor:
    """fuse = LN, Lin(D,4D), LeakyReLU, Dropout, Lin(4D,4D), LeakyReLU, Dropout, Lin(4D,D)
    (models/fused.py:222-229); indices 0,1,4,7 carry parameters."""
    D = x.shape[-1]
    h = F.layer_norm(x, (D,), p[prefix + "0.weight"], p[prefix + "0.bias"])
    h = F.leaky_relu(_lin(p, prefix + "1", h))
    h = F.leaky_relu(_lin(p, prefix + "4", h))
    return _lin(p, prefix + "7", h)


def tabgnn_fused(p: Params, x: Tensor, edge_index: Tensor, edge_attr: Tensor,
                 target_edge_index: Tensor, target_edge_attr: Tensor, num_layers: int,
                 node_dim: int, lp: bool = False, nhead: int = 8, reverse_mp: bool = False,
                 training: bool = True):
    """TABGNNFused.forward, src/nn/models/fused.py:144-175, layer :248-269."""
    B, Ne = target_edge_attr.shape[0], edge_index.shape[1]
    C = p["cls_embedding"].shape[0]
    x_gnn = _lin(p, "node_emb", x.reshape(-1, node_dim))             # :156
    cls = p["cls_embedding"]
    tgt = torch.cat([cls.expand(B, 1, -1), target_edge_attr], dim=1)
    tgt = tab_layer(p, "", tgt, nhead)                               # :160 (no residual)
    ea = torch.cat([cls.expand(Ne, 1, -1), edge_attr], dim=1)
    ea = (ea + tab_layer(p, "", ea, nhead)) / 2                      # :164
    e = _lin(p, "edge_emb", ea.reshape(Ne, -1))                      # :165-166
    x_tab = tgt
    for i in range(num_layers):
        pre = f"backbone.{i}."
        H = x_gnn.shape[1]
        x_tab = x_tab + tab_layer(p, pre, x_tab, nhead) / 2          # :249
        x_cls, x_feat = x_tab[:, 0, :], x_tab[:, 1:, :]
        x_gnn, e = _pna_layer(p, pre, x_gnn, edge_index, e, reverse_mp, training, "avg")   # :252-254
        if not lp:                                                   # :256-268
            z = torch.cat([x_cls, x_gnn[target_edge_index[0]], x_gnn[target_edge_index[1]]], dim=-1)
            D = z.shape[-1]
            z = (z + F.layer_norm(_fuse_mlp(p, pre + "fuse.", z), (D,), p[pre + "fuse_norm.weight"],
                                  p[pre + "fuse_norm.bias"])) / 2
            x_cls = (x_cls + z[:, :C]) / 2
            x_tab = torch.cat([x_cls.unsqueeze(1), x_feat], dim=1)
            index = target_edge_index.flatten()
            emb = torch.cat([z[:, C:C + H], z[:, C + H:]], dim=0)
            uniq, inv = torch.unique(index, return_inverse=True)
            summed = torch.zeros((uniq.numel(), H), dtype=emb.dtype).index_add_(0, inv, emb)
            pooled = summed / torch.bincount(inv).unsqueeze(1).to(emb.dtype)
            x_gnn = x_gnn.clone()
            x_gnn[uniq] = (x_gnn[uniq] + pooled) / 2
    tgt = (x_tab + tgt) / 2                                          # :172
    tgt = _lin(p, "edge_emb", tgt.reshape(B, -1))
    return x_gnn, e, tgt


# ----------------------------------------------------------------- per-column PNA (f4)
def _column_stacks(p: Params, x, edge_index, edge_attr, num_layers, edge_updates, reverse_mp, training):
    """The column loop shared by CPNA.forward (src/nn/gnn/pna.py:215-231) and CPNATAB.forward (:284-297): the node state
    x runs through the K column stacks one after the other; column c's stack only touches edge_attr[:, c, :]."""
    x = _lin(p, "node_emb", x.reshape(x.shape[0], -1))
    src, dst = edge_index[0], edge_index[1]
    cols = []
    for c in range(edge_attr.shape[1]):
        e = edge_attr[:, c, :]
        for i in range(num_layers):
            h = _conv(p, f"col_convs.{c}.{i}.", "pna", reverse_mp, x, edge_index, e, AGG_PNAS)
            h = PF.batch_norm(sub_inplace(p, f"col_batch_norms.{c}.{i}."), h, training)
            x = (x + F.relu(h)) / 2
            if edge_updates:
                e = e + _emlp(p, f"col_emlps.{c}.{i}.", torch.cat([x[src], x[dst], e], dim=-1)) / 2
        cols.append(e)
    return x, torch.stack(cols, dim=1)


def cpna(p: Params, x, edge_index, edge_attr, num_layers, edge_updates=True, reverse_mp=False, training=True):
    """CPNA.forward, src/nn/gnn/pna.py:215-231."""
    return _column_stacks(p, x, edge_index, edge_attr, num_layers, edge_updates, reverse_mp, training)


def cpnatab(p: Params, x, edge_index, edge_attr, num_layers, edge_updates=True, reverse_mp=False, training=True, nhead=8):
    """CPNATAB.forward, src/nn/gnn/pna.py:284-302 (dropout 0 / eval): column stacks, then num_layers rounds of attention over
    the K column tokens of every edge row, (t + row_norms[i](row_atts[i](t))) / 2."""
    x, t = _column_stacks(p, x, edge_index, edge_attr, num_layers, edge_updates, reverse_mp, training)
    C = t.shape[-1]
    for i in range(num_layers):
        a = transformer_encoder_layer(p, f"row_atts.{i}.", t, nhead)
        t = (t + F.layer_norm(a, (C,), p[f"row_norms.{i}.weight"], p[f"row_norms.{i}.bias"])) / 2
    return x, t


# ----------------------------------------------------------------- self-supervised losses / ranking metric (f3)
def lp_loss(pos_pred, neg_pred):
    """SSLoss.lp_loss, src/utils/loss.py:11-12."""
    return -torch.log(pos_pred + 1e-12).mean() - torch.log(1 - neg_pred + 1e-12).mean()


def mcm_loss(cat_out, num_out, y, num_numerical):
    """SSLoss.mcm_loss, src/utils/loss.py:41-72, sample by sample as the reference loops."""
    cat_loss = torch.zeros((), dtype=num_out.dtype)
    num_loss = torch.zeros((), dtype=num_out.dtype)
    acc = t_c = t_n = 0
    for i in range(y.shape[0]):
        val, idx = y[i, 0], int(y[i, 1])
        if idx >= num_numerical:
            logits = cat_out[idx - num_numerical][i]
            cat_loss = cat_loss + F.cross_entropy(logits.unsqueeze(0), val.long().reshape(1), reduction="sum")
            acc += int(torch.argmax(logits) == val.long())
            t_c += 1
        else:
            num_loss = num_loss + (num_out[i, idx] - val) ** 2
            t_n += 1
    if t_c == 0:
        total = torch.sqrt(num_loss / t_n)
    elif t_n == 0:
        total = cat_loss / t_c
    else:
        total = cat_loss / t_c + torch.sqrt(num_loss / t_n)
    return total, (cat_loss, t_c, acc), (num_loss, t_n)


def mrr(pos_pred, neg_pred, ks, num_neg):
    """SSMetric.mrr, src/utils/metric.py:11-52: rank of each positive in the descending order of [negatives, positive]
    (stable sort: the positive comes last among equal scores); returns (mrr, {hits@k})."""
    import numpy as np
    pos = pos_pred.detach().cpu().numpy().flatten()
    neg = neg_pred.detach().cpu().numpy().flatten().reshape(len(pos), num_neg)
    ranks = []
    for pz, ng in zip(pos, neg):
        comb = np.concatenate([ng, [pz]])
        ranks.append(int((-comb).argsort(kind="stable").argsort(kind="stable")[-1]) + 1)
    ranks = np.asarray(ranks)
    return float(np.mean(1.0 / ranks)), {f"hits@{k}": float(np.mean(ranks <= k)) for k in ks}, ranks


# ----------------------------------------------------------------- read-out heads
def classifier_head(p: Params, x: Tensor, edge_index: Tensor, edge_attr: Tensor) -> Tensor:
    """ClassifierHead.forward (eval / dropout 0), src/nn/gnn/decoder.py:17-21; keys mlp.{0,3,6}."""
    H = x.shape[1]
    h = x[edge_index.T].reshape(-1, 2 * H).relu()
    h = torch.cat((h, edge_attr.reshape(-1, edge_attr.shape[1])), 1)
    h = _lin(p, "mlp.0", h).relu()
    h = _lin(p, "mlp.3", h).relu()
    return _lin(p, "mlp.6", h)


def link_pred_head(p: Params, x, pos_ei, pos_ea, neg_ei, neg_ea):
    """LinkPredHead.forward (eval / dropout 0), src/nn/gnn/decoder.py:62-71; keys mlp.{0,3,6}."""
    H = x.shape[1]

    def score(ei, ea):
        h = torch.cat((x[ei.T].reshape(-1, 2 * H).relu(), ea.reshape(-1, ea.shape[1])), 1)
        h = _lin(p, "mlp.0", h).relu()
        h = _lin(p, "mlp.3", h).relu()
        return torch.sigmoid(_lin(p, "mlp.6", h))
    return score(pos_ei, pos_ea), score(neg_ei, neg_ea)
