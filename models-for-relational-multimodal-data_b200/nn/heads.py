"""Read-out heads of the reference (/root/reference/src/nn/gnn/decoder.py): ``ClassifierHead`` (:5-21),
``NodeClassificationHead`` (:23-33), ``LinkPredHead`` (:34-71), ``LinkPredFusedHead`` (:73-102).  Same constructors,
``forward`` signatures and state_dict keys (``mlp.{0,3,6}.{weight,bias}``).

The edge heads start with ``x[edge_index.T] -> relu -> cat(edge_attr) -> Linear``.  relu is element-wise, so the first
Linear is hoisted to the node rows: ``Rs = relu(x) Ws^T``, ``Rd = relu(x) Wd^T`` (N rows, node linears), ``T = edge_attr We^T + b``
(E rows), and one kernel computes ``relu(Rs[src] + Rd[dst] + T)`` (csrc/readout.cu) -- no [E, 3H] concatenation, the gathers
move rows of the output width, and the backward sums the gradient per node over the CSC / CSR orders without atomics.
The first layer's output width is padded with zero rows of the weight to the next kernel width (50 -> 64); the zero
columns stay zero through relu / dropout and meet zero-padded input columns of the second Linear.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F
from torch import Tensor, nn

from .. import functional as BF


def _mlp(sizes, dropout):
    layers = []
    for i in range(len(sizes) - 1):
        layers.append(nn.Linear(sizes[i], sizes[i + 1]))
        if i < len(sizes) - 2:
            layers += [nn.ReLU(), nn.Dropout(dropout)]
    return nn.Sequential(*layers)            # Linear at 0, 3, 6: the reference's key layout


def _pad_rows(W: Tensor, rows: int) -> Tensor:
    return W if W.shape[0] == rows else F.pad(W, (0, 0, 0, rows - W.shape[0]))


def _edge_head_first_layer(mlp: nn.Sequential, H: int, x: Tensor, edge_index: Tensor, edge_attr: Tensor) -> Tensor:
    """relu(Linear_0([relu(x[src]) | relu(x[dst]) | edge_attr])) as [E, Wp] (Wp = padded width), dropout applied by the caller"""
    lin0 = mlp[0]
    O = lin0.weight.shape[0]
    Wp = BF._pad_width(O)
    ea = edge_attr.reshape(-1, edge_attr.shape[1])
    Ws, Wd, We = lin0.weight[:, :H], lin0.weight[:, H:2 * H], lin0.weight[:, 2 * H:]
    xr = BF.relu_dropout(x, 0.0)
    Rs, Rd = BF.linear_group(xr, (_pad_rows(Ws, Wp), _pad_rows(Wd, Wp)), (None, None))
    b = lin0.bias if O == Wp else F.pad(lin0.bias, (0, Wp - O))
    T = BF.linear_group(ea, (_pad_rows(We, Wp),), (b,))[0]
    return BF.pair_gather_act(Rs, Rd, T, edge_index, relu=True)


def _rest(mlp: nn.Sequential, h: Tensor) -> Tensor:
    """layers 2.. of the head on the padded first-layer output (zero-padded input columns of Linear_3)"""
    h = mlp[2](h)                                            # Dropout
    lin3 = mlp[3]
    W3 = lin3.weight if lin3.weight.shape[1] == h.shape[1] else F.pad(lin3.weight, (0, h.shape[1] - lin3.weight.shape[1]))
    h = F.linear(h, W3, lin3.bias)
    for m in list(mlp)[4:]:
        h = m(h)
    return h


class ClassifierHead(nn.Module):
    def __init__(self, n_classes=1, n_hidden=128, dropout=0.5, e_hidden=None):
        super().__init__()
        self.n_hidden = n_hidden
        self.e_hidden = n_hidden if e_hidden is None else e_hidden
        self.mlp = _mlp([n_hidden * 2 + self.e_hidden, 50, 25, n_classes], dropout)

    def forward(self, x, edge_index, edge_attr):
        return _rest(self.mlp, _edge_head_first_layer(self.mlp, self.n_hidden, x, edge_index, edge_attr))


class NodeClassificationHead(nn.Module):
    def __init__(self, n_classes=1, n_hidden=128, dropout=0.5):
        super().__init__()
        self.n_hidden = n_hidden
        self.mlp = _mlp([n_hidden, 50, 25, n_classes], dropout)

    def forward(self, x):
        return self.mlp(x)


class _XavierMixin:
    def reset_parameters(self) -> None:
        for p in self.mlp.parameters():
            if p.dim() > 1:
                torch.nn.init.xavier_uniform_(p)


class LinkPredHead(_XavierMixin, nn.Module):
    def __init__(self, n_classes=1, n_hidden=128, dropout=0.5) -> None:
        super().__init__()
        self.n_hidden, self.n_classes, self.dropout = n_hidden, n_classes, dropout
        self.mlp = _mlp([n_hidden * 3, n_hidden, 25, n_classes], dropout)
        self.reset_parameters()

    def forward(self, x, pos_edge_index, pos_edge_attr, neg_edge_index, neg_edge_attr):
        # positives and negatives share x and the weights: one pass over the concatenated edge list
        # (the node linears Rs / Rd are computed once instead of twice)
        nb = pos_edge_index.shape[1]
        ei = torch.cat([pos_edge_index, neg_edge_index], dim=1)
        ea = torch.cat([pos_edge_attr.reshape(-1, pos_edge_attr.shape[1]), neg_edge_attr.reshape(-1, neg_edge_attr.shape[1])], dim=0)
        out = torch.sigmoid(_rest(self.mlp, _edge_head_first_layer(self.mlp, self.n_hidden, x, ei, ea)))
        return out[:nb], out[nb:]


class LinkPredFusedHead(_XavierMixin, nn.Module):
    def __init__(self, n_classes=1, n_hidden=128, dropout=0.5) -> None:
        super().__init__()
        self.n_hidden, self.n_classes, self.dropout = n_hidden, n_classes, dropout
        self.mlp = _mlp([n_hidden * 3, n_hidden, 25, n_classes], dropout)
        self.reset_parameters()

    def forward(self, pos_emb, neg_emb):
        nb = pos_emb.shape[0]
        out = torch.sigmoid(self.mlp(torch.cat([pos_emb, neg_emb], dim=0)))
        return out[:nb], out[nb:]
