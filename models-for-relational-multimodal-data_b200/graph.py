"""Graph structure on the device: CSC (destination-sorted) and, for reverse message passing,
CSR (source-sorted) segment indices, degree histogram and batch relabel.

Replaces torch_geometric ``to_csc`` / ``NeighborSampler.edge_permutation`` as reached from
/root/reference/src/datasets/util/graph.py:31-57 and ibm_transactions_for_aml.py:91,
``degree`` + bincount at main.py:284-286 / utils.py:179-182, and the relabel loops of
src/utils/batch_processing.py:65-72.  All results are bit-exact (stable sort).
"""
from __future__ import annotations

import ctypes as C
from collections import OrderedDict
from dataclasses import dataclass, field
from typing import Optional, Tuple

import torch

from . import _lib as L

import os as _os
LIGHT_MAX = int(_os.environ.get("B200MP_LIGHT_MAX", 512))     # segments longer than this are reduced by the chunked (multi-CTA) path
CHUNK = int(_os.environ.get("B200MP_CHUNK", 1024))            # rows per chunk of a heavy segment


@dataclass
class SegmentIndex:
    """Edges sorted stably by the node they are aggregated at (``key``)."""
    num_nodes: int
    num_edges: int
    perm: torch.Tensor       # int64 [E]   position in the input edge list of sorted edge k
    ptr: torch.Tensor        # int64 [N+1] segment boundaries (CSC colptr)
    other: torch.Tensor      # int64 [E]   other endpoint of sorted edge k (CSC row)
    nbr32: torch.Tensor      # int32 [E]   = other
    seg32: torch.Tensor      # int32 [E]   aggregation node of sorted edge k
    eid32: Optional[torch.Tensor] = None   # int32 [E] row of edge k in e-storage (None = k)
    _heavy: Optional[tuple] = field(default=None, repr=False)
    _perm32: Optional[torch.Tensor] = field(default=None, repr=False)
    _by_nbr: Optional[dict] = field(default=None, repr=False)

    def by_neighbour(self, num_rows: int) -> "SegmentIndex":
        """The sorted edges grouped (stably) by their OTHER end point: ``perm`` = positions in this index's order, ``ptr`` over
        ``num_rows`` neighbour ids.  Used by the atomic-free backward (functional.set_deterministic): gradient rows written in
        this index's order are summed per neighbour row without scatters.  Built once per structure and row count."""
        if self._by_nbr is None:
            self._by_nbr = {}
        if num_rows not in self._by_nbr:
            self._by_nbr[num_rows] = build_segment_index(self.other, self.other, int(num_rows))
        return self._by_nbr[num_rows]

    def perm32(self):
        """perm as int32 (row map for kernels that read / write the caller's edge order)."""
        if getattr(self, "_perm32", None) is None:
            self._perm32 = self.perm.to(torch.int32).contiguous()
        return self._perm32

    def heavy(self):
        """(n_heavy, heavy_node32, heavy_chunk_ptr32, n_chunks, chunk_heavy32, chunk_begin64):
        work list of the chunked reduction for segments longer than LIGHT_MAX (built once)."""
        if self._heavy is None:
            deg = self.ptr[1:] - self.ptr[:-1]
            hv = (deg > LIGHT_MAX).nonzero().flatten()
            nh = int(hv.numel())
            if nh == 0:
                self._heavy = (0, None, None, 0, None, None)
            else:
                nch = (deg[hv] + CHUNK - 1) // CHUNK
                cptr = torch.zeros(nh + 1, dtype=torch.int64, device=deg.device)
                cptr[1:] = torch.cumsum(nch, 0)
                nc = int(cptr[-1])
                chunk_heavy = torch.repeat_interleave(torch.arange(nh, device=deg.device), nch)
                within = torch.arange(nc, device=deg.device) - cptr[chunk_heavy]
                begin = (self.ptr[hv][chunk_heavy] + within * CHUNK).contiguous()
                self._heavy = (nh, hv.to(torch.int32).contiguous(), cptr.to(torch.int32).contiguous(), nc,
                               chunk_heavy.to(torch.int32).contiguous(), begin)
        return self._heavy


def build_segment_index(key: torch.Tensor, other_end: torch.Tensor, num_nodes: int) -> SegmentIndex:
    """Stable sort of the edges by ``key`` (b200mp_csc_build)."""
    L.require_cuda(key, other_end)
    key = key.contiguous()
    other_end = other_end.contiguous()
    if key.dtype != torch.int64 or other_end.dtype != torch.int64:
        raise TypeError("edge_index must be int64")
    E, N, dev = int(key.numel()), int(num_nodes), key.device
    lib = L.lib()
    nbytes = C.c_size_t(0)
    L.check(lib.b200mp_csc_workspace_bytes(E, N, C.byref(nbytes)), "b200mp_csc_workspace_bytes")
    ws = torch.empty(nbytes.value, dtype=torch.uint8, device=dev)
    perm = torch.empty(E, dtype=torch.int64, device=dev)
    ptr = torch.empty(N + 1, dtype=torch.int64, device=dev)
    other = torch.empty(E, dtype=torch.int64, device=dev)
    nbr32 = torch.empty(E, dtype=torch.int32, device=dev)
    seg32 = torch.empty(E, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        L.check(lib.b200mp_csc_build(L.ptr(key), L.ptr(other_end), E, N, L.ptr(perm), L.ptr(ptr), L.ptr(other),
                                     L.ptr(nbr32), L.ptr(seg32), L.ptr(ws), nbytes.value, L.stream()),
                "b200mp_csc_build")
    return SegmentIndex(N, E, perm, ptr, other, nbr32, seg32)


def to_csc(edge_index: torch.Tensor, num_nodes: int) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """(colptr, row, perm) as PyG's ``to_csc`` returns them for an edge_index-only graph."""
    s = build_segment_index(edge_index[1], edge_index[0], num_nodes)
    return s.ptr, s.other, s.perm


@dataclass
class EdgeStructure:
    """What the kernels need for one edge_index: e-storage is in ``fwd`` (CSC) order."""
    num_nodes: int
    num_edges: int
    fwd: SegmentIndex                     # sorted by dst; nbr32 = src, seg32 = dst
    rev: Optional[SegmentIndex] = None    # sorted by src; nbr32 = dst, seg32 = src, eid32 -> fwd rows

    @property
    def src32(self):
        return self.fwd.nbr32

    @property
    def dst32(self):
        return self.fwd.seg32

    def ensure_reverse(self, edge_index: torch.Tensor):
        if self.rev is None:
            rev = build_segment_index(edge_index[0], edge_index[1], self.num_nodes)
            inv = torch.empty_like(self.fwd.perm)
            inv[self.fwd.perm] = torch.arange(self.num_edges, device=inv.device)
            rev.eid32 = inv[rev.perm].to(torch.int32).contiguous()
            self.rev = rev
        return self.rev


_CACHE: "OrderedDict[tuple, tuple]" = OrderedDict()
_CACHE_SIZE = 2


def edge_structure(edge_index: torch.Tensor, num_nodes: int, reverse: bool = False) -> EdgeStructure:
    """Build (or fetch) the structure of ``edge_index``.  The cache holds a reference to the
    keyed tensor, so its memory cannot be recycled under a stale entry; in-place edits bump
    ``_version`` and miss."""
    if edge_index.dim() != 2 or edge_index.shape[0] != 2:
        raise ValueError("edge_index must have shape [2, E]")
    key = (edge_index.data_ptr(), tuple(edge_index.shape), tuple(edge_index.stride()), edge_index._version,
           int(num_nodes), str(edge_index.device))
    hit = _CACHE.get(key)
    if hit is not None:
        _CACHE.move_to_end(key)
        st = hit[1]
    else:
        fwd = build_segment_index(edge_index[1], edge_index[0], num_nodes)
        st = EdgeStructure(int(num_nodes), int(edge_index.shape[1]), fwd)
        _CACHE[key] = (edge_index, st)
        while len(_CACHE) > _CACHE_SIZE:
            _CACHE.popitem(last=False)
    if reverse:
        st.ensure_reverse(edge_index)
    return st


def clear_cache():
    _CACHE.clear()


def degree(index: torch.Tensor, num_nodes: Optional[int] = None, dtype=None) -> torch.Tensor:
    """torch_geometric.utils.degree (main.py:286): occurrences of each node id in ``index``."""
    L.require_cuda(index)
    N = int(index.max()) + 1 if num_nodes is None else int(num_nodes)
    out = torch.empty(N, dtype=torch.int64, device=index.device)
    with torch.cuda.device(index.device):
        L.check(L.lib().b200mp_bincount(L.ptr(index.contiguous()), index.numel(), L.ptr(out), N, L.stream()),
                "b200mp_bincount")
    return out.to(dtype if dtype is not None else torch.float32)


def degree_histogram(ptr: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """(deg[N], hist[max_deg+1]) int64 from a CSC pointer (main.py:284-286 + utils.py:179-182)."""
    L.require_cuda(ptr)
    N = ptr.numel() - 1
    deg = torch.empty(N, dtype=torch.int64, device=ptr.device)
    mx = torch.zeros(1, dtype=torch.int64, device=ptr.device)
    with torch.cuda.device(ptr.device):
        L.check(L.lib().b200mp_degree(L.ptr(ptr), N, L.ptr(deg), L.ptr(mx), L.stream()), "b200mp_degree")
        nb = int(mx.item()) + 1
        hist = torch.empty(nb, dtype=torch.int64, device=ptr.device)
        L.check(L.lib().b200mp_bincount(L.ptr(deg), N, L.ptr(hist), nb, L.stream()), "b200mp_bincount")
    return deg, hist


def relabel(ids: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """(nodes, local): nodes = unique(ids) ascending, local[i] = rank of ids[i]
    (batch_processing.py:65-72)."""
    L.require_cuda(ids)
    ids = ids.contiguous().view(-1)
    M, dev = ids.numel(), ids.device
    nbytes = C.c_size_t(0)
    L.check(L.lib().b200mp_relabel_workspace_bytes(M, C.byref(nbytes)), "b200mp_relabel_workspace_bytes")
    ws = torch.empty(nbytes.value, dtype=torch.uint8, device=dev)
    nodes = torch.empty(M, dtype=torch.int64, device=dev)
    local = torch.empty(M, dtype=torch.int64, device=dev)
    nu = torch.zeros(1, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        L.check(L.lib().b200mp_relabel(L.ptr(ids), M, L.ptr(nodes), L.ptr(nu), L.ptr(local), L.ptr(ws), nbytes.value,
                                       L.stream()), "b200mp_relabel")
    return nodes[: int(nu.item())], local
