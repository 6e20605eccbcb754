"""-m gpu: b200mp_relabel_static -- the batch relabel with its workspace in a file-scope `static __device__` array inside the
library (no cudaMalloc on either side of the call; include/b200mp.h has the contract) -- against b200mp_relabel, through the
C ABI, on calls in a row that share the array.  Integer work: bit-exact.

Collected last on purpose and self-contained: it passes alone in a fresh process and after the
other tests.  No reference counterpart."""
import ctypes as C

import pytest
import torch

from b200mp import _lib as L

pytestmark = pytest.mark.gpu
DEV = "cuda"
MAX_IDS = 262144          # B200MP_RELABEL_STATIC_MAX_IDS
ELIMIT = -3               # B200MP_ELIMIT


def _relabel_workspace(ids):
    """the existing entry point: caller-allocated workspace"""
    M = ids.numel()
    nbytes = C.c_size_t(0)
    assert L.lib().b200mp_relabel_workspace_bytes(M, C.byref(nbytes)) == 0
    ws = torch.empty(nbytes.value, dtype=torch.uint8, device=DEV)
    nodes = torch.full((max(M, 1),), -7, dtype=torch.int64, device=DEV)
    local = torch.full((max(M, 1),), -7, dtype=torch.int64, device=DEV)
    nu = torch.full((1,), -7, dtype=torch.int64, device=DEV)
    rc = L.lib().b200mp_relabel(L.ptr(ids), M, L.ptr(nodes), L.ptr(nu), L.ptr(local), L.ptr(ws), nbytes.value, L.stream())
    assert rc == 0, rc
    torch.cuda.synchronize()
    return nodes, int(nu.item()), local


def _relabel_static(ids, nodes, nu, local):
    """the new entry point: outputs preallocated by the caller, nothing allocated inside the call"""
    rc = L.lib().b200mp_relabel_static(L.ptr(ids), ids.numel(), L.ptr(nodes), L.ptr(nu), L.ptr(local), L.stream())
    torch.cuda.synchronize()
    return rc


def _ids(M, hi, seed):
    g = torch.Generator(device=DEV).manual_seed(seed)
    return torch.randint(0, hi, (M,), device=DEV, generator=g, dtype=torch.int64) * 7 - 3 * hi      # negative ids too


def test_relabel_static_matches_relabel_on_calls_in_a_row():
    nodes = torch.empty(MAX_IDS, dtype=torch.int64, device=DEV)
    local = torch.empty(MAX_IDS, dtype=torch.int64, device=DEV)
    nu = torch.empty(1, dtype=torch.int64, device=DEV)
    # the largest accepted size first, then smaller ones: every later call finds the earlier call's contents in the shared array
    cases = [(MAX_IDS, 50_000, 1), (100_001, 1 << 40, 2), (4097, 13, 3), (1, 5, 4), (100_001, 1 << 40, 2), (MAX_IDS, 3, 5)]
    for M, hi, seed in cases:
        ids = _ids(M, hi, seed)
        want_nodes, want_n, want_local = _relabel_workspace(ids)
        nodes.fill_(-9), local.fill_(-9), nu.fill_(-9)
        before = torch.cuda.memory_allocated()
        rc = _relabel_static(ids, nodes, nu, local)
        assert rc == 0, (M, rc)
        assert torch.cuda.memory_allocated() == before            # torch's allocator was not asked for anything either
        n = int(nu.item())
        assert n == want_n, (M, n, want_n)
        assert torch.equal(nodes[:n], want_nodes[:n]), M
        assert torch.equal(local[:M], want_local[:M]), M
        # and against the definition itself
        u, inv = torch.unique(ids, return_inverse=True)
        assert torch.equal(nodes[:n], u) and torch.equal(local[:M], inv)


def test_relabel_static_limits_and_empty():
    nodes = torch.full((8,), -9, dtype=torch.int64, device=DEV)
    local = torch.full((8,), -9, dtype=torch.int64, device=DEV)
    nu = torch.full((1,), -9, dtype=torch.int64, device=DEV)
    ids = torch.zeros(8, dtype=torch.int64, device=DEV)
    # beyond the maximum: refused before anything is launched, outputs untouched
    rc = L.lib().b200mp_relabel_static(L.ptr(ids), MAX_IDS + 1, L.ptr(nodes), L.ptr(nu), L.ptr(local), L.stream())
    torch.cuda.synchronize()
    assert rc == ELIMIT
    assert int(nu.item()) == -9 and bool((nodes == -9).all()) and bool((local == -9).all())
    with pytest.raises(L.B200MPError, match="fixed maximum"):
        L.check(rc, "b200mp_relabel_static")
    assert L.lib().b200mp_relabel_static(L.ptr(ids), -1, L.ptr(nodes), L.ptr(nu), L.ptr(local), L.stream()) == -1
    # empty input: n_unique = 0, like b200mp_relabel
    assert _relabel_static(ids[:0], nodes, nu, local) == 0
    assert int(nu.item()) == 0
    # and the array still serves a normal call afterwards
    ids = torch.tensor([5, 3, 5, 9, 3, 3, -1, 9], dtype=torch.int64, device=DEV)
    assert _relabel_static(ids, nodes, nu, local) == 0
    assert int(nu.item()) == 4
    assert nodes[:4].tolist() == [-1, 3, 5, 9] and local.tolist() == [2, 1, 2, 3, 1, 1, 0, 3]
