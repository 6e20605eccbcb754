"""-m gpu: the device neighbour sampler against the oracle's take-all
restatement of PyG NeighborSampler / pyg-lib neighbor_sample (oracle/index_ops.py, App. A.3):
bit-exact positions, node order and (row, col, idx) tables in the take-all regime; for hubs beyond the
fan-out: count, range, distinctness, and uniform coverage of the slice."""
import numpy as np
import pytest
import torch

import b200mp
from oracle import index_ops as OI

pytestmark = pytest.mark.gpu


def graph(n, e, seed, hub=None):
    g = torch.Generator().manual_seed(seed)
    ei = torch.randint(0, n, (2, e), generator=g)
    if hub:
        ei[1, :hub] = 3                     # node 3 gets `hub` in-edges
    return ei


@pytest.mark.parametrize("n,e,b,fan", [(3000, 9000, 50, (100, 100)), (500, 4000, 20, (100, 100)), (200, 300, 7, (-1, -1))])
def test_take_all_bit_exact(n, e, b, fan):
    ei = graph(n, e, n + e)
    perm, colptr, row = OI.csc_build(ei.numpy(), n)
    sampler = b200mp.NeighborSampler(ei.cuda(), n, fan)
    assert np.array_equal(sampler.edge_permutation.cpu().numpy(), perm)
    ids = torch.randperm(e, generator=torch.Generator().manual_seed(1))[:b]
    seeds = torch.stack([ei[0, ids], ei[1, ids], ids], 1)
    out_edge, nodes = sampler.sample_from_edges(seeds[:, 0].cuda(), seeds[:, 1].cuda())
    ref_edge, ref_nodes = OI.sample_from_edges_take_all(colptr, row, seeds[:, 0].numpy(), seeds[:, 1].numpy(), fan)
    assert np.array_equal(out_edge.cpu().numpy(), ref_edge)
    assert np.array_equal(nodes.cpu().numpy(), ref_nodes)
    table = ei.t().contiguous()
    got = sampler.sample_neighbors(seeds.cuda(), table.cuda())
    ref = OI.sample_neighbors(table.numpy(), perm, colptr, row, seeds.numpy(), fan)
    for a, r in zip(got, ref):
        assert np.array_equal(a.cpu().numpy(), r)
    # the scratch state is restored: a second call gives the same answer
    again, _ = sampler.sample_from_edges(seeds[:, 0].cuda(), seeds[:, 1].cuda())
    assert torch.equal(again, out_edge)


def test_hub_subsampling_valid_and_uniform():
    n, e, hub, fan = 1000, 6000, 2500, 100
    ei = graph(n, e, 5, hub=hub)
    perm, colptr, row = OI.csc_build(ei.numpy(), n)
    sampler = b200mp.NeighborSampler(ei.cuda(), n, (fan,))
    lo, hi = int(colptr[3]), int(colptr[4])
    assert hi - lo >= hub
    hits = np.zeros(hi - lo)
    for s in range(200):
        out_edge, _ = sampler.sample_from_edges(torch.tensor([3]).cuda(), torch.tensor([3]).cuda(), seed=s)
        p = out_edge.cpu().numpy()
        assert p.size == fan and len(set(p.tolist())) == fan and p.min() >= lo and p.max() < hi
        hits[p - lo] += 1
    exp = 200 * fan / (hi - lo)
    z = (((hits - exp) ** 2 / exp).sum() - (hi - lo - 1)) / np.sqrt(2 * (hi - lo - 1))
    print(f"hub slice {hi - lo}, chi-square z {z:.2f}")
    assert abs(z) < 5
