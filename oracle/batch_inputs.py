"""ORACLE (test infrastructure, never on the product path): numpy restatement of the reference's batch composition,
/root/reference/src/utils/batch_processing.py -- ``mcm_inputs`` (:83-104), ``lp_inputs`` (:105-150), ``addEgoIDs`` (:153-160) --
on top of oracle/index_ops.sample_neighbors (take-all regime).  The negative edges themselves are random in the reference
(std::random_device); they are judged by index_ops.negative_samples_valid, everything else is compared bit for bit.

Tables are dicts name -> ndarray [rows, ...] (the feat_dict of a TensorFrame)."""
import numpy as np

from . import index_ops as OI


def _subgraph(edges_table, perm, colptr, row, feats, seed_edges, num_neighbors):
    s_row, s_col, s_idx = OI.sample_neighbors(edges_table, perm, colptr, row, seed_edges, num_neighbors)     # dataset.sample_neighbors
    edge_attr = {k: v[s_idx] for k, v in feats.items()}                       # tensor_frame.__getitem__(idx)
    nodes = np.unique(np.concatenate([s_row, s_col]))                         # :112 (sorted)
    n_id_map = {int(v): i for i, v in enumerate(nodes)}                       # :116
    local = np.asarray([n_id_map[int(v)] for v in np.concatenate([s_row, s_col])], dtype=np.int64)     # :117-120
    edge_index = local.reshape(2, -1)
    node_feats = np.ones((nodes.size, 1), dtype=np.float32)                   # :114
    return node_feats, edge_index, edge_attr, s_idx


def add_ego_ids(x, seed_edge_index):
    ids = np.zeros((x.shape[0], 1), dtype=x.dtype)
    ids[np.unique(seed_edge_index.reshape(-1))] = 1
    return np.concatenate([x, ids], axis=1)


def mcm_inputs(edges_table, perm, colptr, row, feats, seed_edges, num_neighbors=(100, 100), ego=False):
    B = seed_edges.shape[0]
    node_feats, edge_index, edge_attr, _ = _subgraph(edges_table, perm, colptr, row, feats, seed_edges, num_neighbors)
    if ego:
        node_feats = add_ego_ids(node_feats, edge_index[:, :B])
    return node_feats, edge_index, edge_attr, edge_index[:, :B], {k: v[:B] for k, v in edge_attr.items()}


def lp_inputs(edges_table, perm, colptr, row, feats, seed_edges, num_neg_samples=64, num_neighbors=(100, 100), ego=False):
    """everything lp_inputs returns except the negative edges (target_edge_index[:, B:])"""
    B = seed_edges.shape[0]
    node_feats, edge_index, edge_attr, _ = _subgraph(edges_table, perm, colptr, row, feats, seed_edges, num_neighbors)
    if ego:
        node_feats = add_ego_ids(node_feats, edge_index[:, :B])
    neigh_edge_index = edge_index[:, B:]
    neigh_edge_attr = {k: v[B:] for k, v in edge_attr.items()}
    pos_edge_index = edge_index[:, :B]
    target = {}
    for k, v in edge_attr.items():                                            # :131-142: each positive row num_neg times, adjacent
        pos = v[:B]
        rep = [np.repeat(r[None], num_neg_samples, axis=0) for r in pos]
        target[k] = np.concatenate([pos] + rep, axis=0) if B else pos
    return node_feats, edge_index, edge_attr, neigh_edge_index, neigh_edge_attr, pos_edge_index, target
