"""Batch construction on the device: the reference's ``graph_inputs`` / ``mcm_inputs`` / ``lp_inputs`` /
``addEgoIDs`` (/root/reference/src/utils/batch_processing.py:58-160) composed from the device sampler
(``NeighborSampler.sample_neighbors``, ibm_transactions_for_aml.py:61-112), the relabel kernel (``unique`` + rank instead
of the Python dict / ``np.vectorize`` loops at :65-72, :87-97, :112-121), TensorFrame row gathers, the x``num_neg``
attribute replication (:133-143, a per-row Python loop in the reference) as ONE index gather per stype, and the device
negative sampler (:145; the reference converts the batch to Python lists and calls its C++ module).  Nothing leaves
the device; the only host reads are the sizes the sampler needs.

``BatchBuilder`` plays the role of the reference's dataset object in those functions (``dataset.sample_neighbors``,
``dataset.tensor_frame``).
"""
from __future__ import annotations

from typing import Optional

import torch

from .graph import relabel
from .sampling import NeighborSampler, generate_negative_samples


def addEgoIDs(x: torch.Tensor, seed_edge_index: torch.Tensor) -> torch.Tensor:
    """batch_processing.py:153-160: one extra column, 1 for the end points of the seed edges"""
    ids = torch.zeros((x.shape[0], 1), device=x.device, dtype=x.dtype)
    ids[seed_edge_index.reshape(-1)] = 1
    return torch.cat([x, ids], dim=1)


def repeat_rows(tf, times: int):
    """every row of a TensorFrame ``times`` times, copies adjacent (the replication loop of batch_processing.py:133-143)"""
    from .nn.encoders import TensorFrame
    n = tf.num_rows
    idx = torch.arange(n, device=tf.device).repeat_interleave(int(times))
    return TensorFrame({k: v[idx] for k, v in tf.feat_dict.items()}, tf.col_names_dict, None)


def concat_rows(a, b):
    from .nn.encoders import TensorFrame
    return TensorFrame({k: torch.cat([a.feat_dict[k], b.feat_dict[k]], dim=0) for k in a.feat_dict}, a.col_names_dict, None)


class BatchBuilder:
    """edge_index [2, E] (global account ids, row e = transaction e), tensor_frame: the edge table (row e = transaction e),
    both on the device; num_neighbors as the reference's ``khop_neighbors`` ([100, 100])."""

    def __init__(self, edge_index: torch.Tensor, num_nodes: int, tensor_frame, num_neighbors=(100, 100)):
        self.sampler = NeighborSampler(edge_index, num_nodes, num_neighbors)
        self.edge_table = edge_index.t().contiguous()
        self.tensor_frame = tensor_frame

    def sample_neighbors(self, edges: torch.Tensor, seed: Optional[int] = None):
        return self.sampler.sample_neighbors(edges, self.edge_table, seed)

    def _subgraph(self, edges, seed):
        row, col, idx = self.sample_neighbors(edges, seed)
        edge_attr = self.tensor_frame[idx]
        nodes, local = relabel(torch.cat([row, col]))                 # nodes ascending, local = rank (the n_id_map)
        edge_index = local.view(2, -1)
        node_feats = torch.ones((nodes.numel(), 1), device=edge_index.device)
        return node_feats, edge_index, edge_attr, nodes, idx

    def graph_inputs(self, edges: torch.Tensor, y: torch.Tensor, ego: bool = False, seed: Optional[int] = None):
        """batch_processing.py:58-81 -> (node_feats, edge_index, edge_attr, y); edges = batch.y[:, -3:] = (src, dst, id)"""
        node_feats, edge_index, edge_attr, _, _ = self._subgraph(edges, seed)
        if ego:
            node_feats = addEgoIDs(node_feats, edge_index[:, :edges.shape[0]])
        return node_feats, edge_index, edge_attr, y.to(torch.long)

    def mcm_inputs(self, edges: torch.Tensor, ego: bool = False, seed: Optional[int] = None):
        """batch_processing.py:83-104 -> (node_feats, edge_index, edge_attr, target_edge_index, target_edge_attr)"""
        B = edges.shape[0]
        node_feats, edge_index, edge_attr, _, _ = self._subgraph(edges, seed)
        if ego:
            node_feats = addEgoIDs(node_feats, edge_index[:, :B])
        return node_feats, edge_index, edge_attr, edge_index[:, :B], edge_attr[torch.arange(B, device=edge_index.device)]

    def lp_inputs(self, edges: torch.Tensor, num_neg_samples: int = 64, ego: bool = False, seed: Optional[int] = None):
        """batch_processing.py:105-150 -> (node_feats, edge_index, edge_attr, neigh_edge_index, neigh_edge_attr,
        target_edge_index, target_edge_attr): positives first, then per positive num_neg/2 (src, n) and num_neg/2 (n, dst)
        negatives; target_edge_attr = the B positive rows followed by each of them repeated num_neg times"""
        B = edges.shape[0]
        node_feats, edge_index, edge_attr, _, _ = self._subgraph(edges, seed)
        if ego:
            node_feats = addEgoIDs(node_feats, edge_index[:, :B])
        dev = edge_index.device
        rest = torch.arange(B, edge_index.shape[1], device=dev)
        neigh_edge_index, neigh_edge_attr = edge_index[:, B:], edge_attr[rest]
        pos_edge_index, pos_edge_attr = edge_index[:, :B], edge_attr[torch.arange(B, device=dev)]
        target_edge_attr = concat_rows(pos_edge_attr, repeat_rows(pos_edge_attr, num_neg_samples))
        neg_edge_index = generate_negative_samples(edge_index.contiguous(), pos_edge_index.contiguous(), num_neg_samples, seed=seed,
                                                   num_nodes=node_feats.shape[0])
        target_edge_index = torch.cat([pos_edge_index, neg_edge_index], dim=1)
        return node_feats, edge_index, edge_attr, neigh_edge_index, neigh_edge_attr, target_edge_index, target_edge_attr
