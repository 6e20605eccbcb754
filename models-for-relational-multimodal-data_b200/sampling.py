"""Batch construction on the device.

``generate_negative_samples`` is the drop-in for the reference's pybind11 module
``src.primitives.negative_sampling`` (/root/reference/src/primitives/negative_sampling/negative_sampling.cpp:10-76,
called at src/utils/batch_processing.py:145 with Python lists): same arguments, same output layout
(``[2][B * num_neg]``: per positive, num_neg/2 edges (src, n) then num_neg/2 edges (n, dst)), same
``ValueError`` for ``num_neg_samples <= 0``.  It also takes tensors (no host round trip) and a seed: the
reference draws from ``std::random_device`` and cannot be reproduced, this sampler can.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib as L

_seed_counter = [0x5eed]


def generate_negative_samples(edge_index, pos_edge_index, num_neg_samples: int, seed: Optional[int] = None,
                              num_nodes: Optional[int] = None, device=None):
    if num_neg_samples <= 0:
        raise ValueError("num_neg_samples must be greater than 0")     # negative_sampling.cpp:13-15
    as_list = not torch.is_tensor(edge_index)
    if as_list:
        if not torch.cuda.is_available():
            raise L.B200MPError("b200mp.generate_negative_samples runs on a CUDA device (no CPU fallback)")
        device = torch.device(device if device is not None else "cuda")
        ei = torch.tensor(edge_index, dtype=torch.long).reshape(2, -1).to(device)
        pos = torch.tensor(pos_edge_index, dtype=torch.long).reshape(2, -1).to(device)
    else:
        ei, pos = edge_index.to(torch.long).contiguous(), pos_edge_index.to(torch.long).contiguous()
        L.require_cuda(ei, pos)
        pos = pos.to(ei.device)
    E, B = ei.shape[1], pos.shape[1]
    half = num_neg_samples // 2
    neg = torch.empty((2, B * 2 * half), dtype=torch.long, device=ei.device)
    if B > 0 and half > 0:
        if seed is None:
            _seed_counter[0] += 1
            seed = _seed_counter[0] * 0x9E3779B97F4A7C15 + torch.initial_seed()
        nbytes = C.c_size_t(0)
        L.check(L.lib().b200mp_negative_sample_workspace_bytes(E, C.byref(nbytes)), "b200mp_negative_sample_workspace_bytes")
        ws = torch.empty(nbytes.value, dtype=torch.uint8, device=ei.device)
        with torch.cuda.device(ei.device):
            L.check(L.lib().b200mp_negative_sample(L.ptr(ei), E, int(num_nodes or 0), L.ptr(pos), B, int(num_neg_samples),
                                                   C.c_uint64(seed & (2 ** 64 - 1)), L.ptr(neg), L.ptr(ws), nbytes.value,
                                                   L.stream()), "b200mp_negative_sample")
    return neg.tolist() if as_list else neg


class NeighborSampler:
    """CSC neighbour sampler on the device: ``sample_from_edges`` mirrors PyG
    ``NeighborSampler.sample_from_edges(EdgeSamplerInput(None, row, col))`` as the reference uses it
    (src/datasets/util/graph.py:31-57 builds it, ibm_transactions_for_aml.py:88-102 calls it and reads
    only ``out.edge`` and ``edge_permutation``): seeds = unique(cat(row, col)) ascending; per hop, the
    frontier nodes in discovery order, each one's in-edge slice whole when deg <= fan-out (bit-exact),
    else ``fanout`` distinct positions; nodes de-duplicated globally; result = CSC positions, hop-major.
    ``sample_neighbors`` is the dataset method built on it (ibm...:61-112, eth...:62-113)."""

    def __init__(self, edge_index: torch.Tensor, num_nodes: int, num_neighbors=(100, 100)):
        from .graph import edge_structure
        L.require_cuda(edge_index)
        self.num_nodes = int(num_nodes)
        self.num_neighbors = [int(f) for f in num_neighbors]
        seg = edge_structure(edge_index, self.num_nodes).fwd
        self.colptr, self.row, self.edge_permutation = seg.ptr, seg.other, seg.perm
        dev = edge_index.device
        self._visited = torch.zeros(self.num_nodes, dtype=torch.uint8, device=dev)
        self._first = torch.full((self.num_nodes,), -1, dtype=torch.int64, device=dev)       # 0xFF..FF
        self._calls = 0

    def sample_from_edges(self, row: torch.Tensor, col: torch.Tensor, seed: Optional[int] = None):
        """-> (out_edge [CSC positions, hop-major], nodes [seeds then discovered nodes, in order])"""
        from .graph import relabel
        L.require_cuda(row, col)
        dev = self.colptr.device
        lib = L.lib()
        self._calls += 1
        seed = (self._calls * 0x9E3779B97F4A7C15 + torch.initial_seed()) if seed is None else seed
        frontier, _ = relabel(torch.cat([row.to(torch.long), col.to(torch.long)]))
        nodes, edges = [frontier], []
        self._visited[frontier] = 1
        with torch.cuda.device(dev):
            for hop, fan in enumerate(self.num_neighbors):
                F = frontier.numel()
                if F == 0:
                    break
                take = torch.empty(F, dtype=torch.int64, device=dev)
                L.check(lib.b200mp_sample_hop_count(L.ptr(self.colptr), L.ptr(frontier), F, fan, L.ptr(take), L.stream()),
                        "b200mp_sample_hop_count")
                ends = torch.cumsum(take, 0)
                offs = ends - take
                K = int(ends[-1].item())
                pos = torch.empty(K, dtype=torch.int64, device=dev)
                L.check(lib.b200mp_sample_hop_fill(L.ptr(self.colptr), L.ptr(frontier), F, fan, L.ptr(offs),
                                                   C.c_uint64((seed + hop) & (2 ** 64 - 1)), L.ptr(pos), L.stream()),
                        "b200mp_sample_hop_fill")
                edges.append(pos)
                cand = self.row[pos]
                flags = torch.empty(K, dtype=torch.uint8, device=dev)
                L.check(lib.b200mp_first_occurrence(L.ptr(cand), K, L.ptr(self._visited), L.ptr(self._first), L.ptr(flags),
                                                    L.stream()), "b200mp_first_occurrence")
                frontier = cand[flags.bool()]
                nodes.append(frontier)
        nodes = torch.cat(nodes)
        self._visited[nodes] = 0                       # scratch back to its resting state
        out_edge = torch.cat(edges) if edges else torch.empty(0, dtype=torch.int64, device=dev)
        return out_edge, nodes

    def sample_neighbors(self, edges: torch.Tensor, edge_table: torch.Tensor, seed: Optional[int] = None):
        """edges: [B,3] = (src, dst, id) seed edges; edge_table: [E,>=2] global (src, dst) per edge id.
        -> (row, col, idx): the seed edges first, in order, then the sampled edges that are not seeds."""
        edges = edges.to(torch.long)
        row, col, idx = edges[:, 0], edges[:, 1], edges[:, 2]
        out_edge, _ = self.sample_from_edges(row, col, seed)
        e_id = self.edge_permutation[out_edge]
        new = e_id[~torch.isin(e_id, idx)]
        if new.numel() > 0:
            row = torch.cat([row, edge_table[new, 0].to(torch.long)])
            col = torch.cat([col, edge_table[new, 1].to(torch.long)])
            idx = torch.cat([idx, new])
        return row, col, idx
