"""ORACLE (test infrastructure, not product code): numpy restatement of the INTEGER side of the
hot path -- CSC/CSR build, degree histogram, neighbour sampling in the take-all regime, batch
relabel and the negative-sampler validity rule.  Bit-exact targets.

Third-party pieces (PyG 2.5.3 ``to_csc`` / ``NeighborSampler`` / pyg-lib ``neighbor_sample``) are
absent from /root/reference; their published behaviour is restated.
"""
from __future__ import annotations

import numpy as np


def csc_build(edge_index: np.ndarray, num_nodes: int):
    """PyG ``to_csc`` as reached from create_graph (src/datasets/util/graph.py:31-57):
    perm = stable argsort by dst, row = src[perm], colptr = index2ptr(dst[perm]).
    ``perm`` is what the reference reads back as ``sampler.edge_permutation``
    (src/datasets/ibm_transactions_for_aml.py:91)."""
    src, dst = edge_index[0], edge_index[1]
    perm = np.argsort(dst, kind="stable").astype(np.int64)
    colptr = np.zeros(num_nodes + 1, dtype=np.int64)
    np.cumsum(np.bincount(dst, minlength=num_nodes), out=colptr[1:])
    return perm, colptr, src[perm].astype(np.int64)


def csr_build(edge_index: np.ndarray, num_nodes: int):
    """Same, keyed by src (needed for the backward scatter and for reverse_mp,
    gine.py:30-33 / pna.py:41-44 which call the conv on edge_index.flipud())."""
    return csc_build(edge_index[::-1], num_nodes)


def degree_hist(dst: np.ndarray, num_nodes: int):
    """main.py:284-286 + utils.py:179-182: in-degree per node, then bincount of the degrees."""
    deg = np.bincount(dst, minlength=num_nodes).astype(np.int64)
    hist = np.bincount(deg, minlength=int(deg.max()) + 1 if deg.size else 1).astype(np.int64)
    return deg, hist


def relabel(khop_src: np.ndarray, khop_dst: np.ndarray):
    """batch_processing.py:65-72 / ibm_transactions_for_aml.py:166-173: nodes = unique(cat) sorted
    ascending, local id = rank."""
    nodes = np.unique(np.concatenate([khop_src, khop_dst]))
    loc = np.searchsorted(nodes, np.concatenate([khop_src, khop_dst])).astype(np.int64)
    return nodes.astype(np.int64), np.stack([loc[: khop_src.size], loc[khop_src.size:]])


def sample_from_edges_take_all(colptr: np.ndarray, row: np.ndarray, seed_src: np.ndarray,
                               seed_dst: np.ndarray, num_neighbors):
    """NeighborSampler.sample_from_edges -> pyg-lib neighbor_sample (non-disjoint, no
    replacement), deterministic when every visited in-degree <= fan-out ("take-all").
    Seeds = unique(cat(row, col)) ascending; hop h expands, in order, the nodes first reached at
    hop h-1; each node's CSC slice is taken whole; newly seen nodes are appended in slice order.
    Returns the CSC positions (``out.edge``), hop-major.  Raises if a slice exceeds the fan-out
    (sub-sampling RNG of pyg-lib is not reproducible)."""
    seeds = np.unique(np.concatenate([seed_src, seed_dst]))
    seen = {int(v): i for i, v in enumerate(seeds)}
    nodes = [int(v) for v in seeds]
    out_edges = []
    begin, end = 0, len(nodes)
    for fan in num_neighbors:
        for i in range(begin, end):
            v = nodes[i]
            lo, hi = int(colptr[v]), int(colptr[v + 1])
            if fan >= 0 and hi - lo > fan:
                raise ValueError("take-all regime violated: degree > fan-out")
            for pos in range(lo, hi):
                out_edges.append(pos)
                u = int(row[pos])
                if u not in seen:
                    seen[u] = len(nodes)
                    nodes.append(u)
        begin, end = end, len(nodes)
    return np.asarray(out_edges, dtype=np.int64), np.asarray(nodes, dtype=np.int64)


def sample_neighbors(edges_table: np.ndarray, perm: np.ndarray, colptr: np.ndarray, row: np.ndarray,
                     seed_edges: np.ndarray, num_neighbors=(100, 100)):
    """IBMTransactionsAML.sample_neighbors, ibm_transactions_for_aml.py:86-112 (take-all regime):
    e_id = perm[out.edge]; drop ids already among the seeds (torch.isin); seeds first, then new
    edges read from the global edge table.  ``seed_edges`` is [B,3] = (src, dst, id)."""
    s_row, s_col, s_idx = seed_edges[:, 0], seed_edges[:, 1], seed_edges[:, 2]
    out_edge, _ = sample_from_edges_take_all(colptr, row, s_row, s_col, num_neighbors)
    e_id = perm[out_edge]
    new = e_id[~np.isin(e_id, s_idx)]
    if new.size:
        s_row = np.concatenate([s_row, edges_table[new, 0]])
        s_col = np.concatenate([s_col, edges_table[new, 1]])
        s_idx = np.concatenate([s_idx, new])
    return s_row.astype(np.int64), s_col.astype(np.int64), s_idx.astype(np.int64)


def negative_samples_valid(edge_index: np.ndarray, pos_edge_index: np.ndarray, neg: np.ndarray,
                           num_neg: int) -> bool:
    """Validity rule of generate_negative_samples (negative_sampling.cpp:35-72): for positive i,
    the first num_neg/2 negatives are (src_i, n), the next num_neg/2 are (n, dst_i), with
    n in [0, |nodeset|) and n not in {src_i, dst_i} U N(src_i) U N(dst_i) (undirected 1-hop)."""
    nodeset = np.unique(edge_index)
    adj = {}
    for s, d in zip(edge_index[0].tolist(), edge_index[1].tolist()):
        adj.setdefault(s, set()).add(d)
        adj.setdefault(d, set()).add(s)
    half = num_neg // 2
    if neg.shape != (2, 2 * half * pos_edge_index.shape[1]):
        return False
    for i, (s, d) in enumerate(zip(pos_edge_index[0].tolist(), pos_edge_index[1].tolist())):
        bad = {s, d} | adj.get(s, set()) | adj.get(d, set())
        blk = neg[:, i * 2 * half:(i + 1) * 2 * half]
        a, b = blk[:, :half], blk[:, half:]
        if not (np.all(a[0] == s) and np.all(b[1] == d)):
            return False
        for n in np.concatenate([a[1], b[0]]).tolist():
            if n in bad or n < 0 or n >= nodeset.size:
                return False
    return True
