"""Negative sampler.  The reference's RNG is seeded from
std::random_device, so parity is: the output constraints of negative_sampling.cpp:10-76 (layout,
range, exclusion of src / dst / their undirected neighbours) and the distribution of the drawn nodes.

not gpu: the oracle port against the COMPILED reference (oracle/_ref, when built) and the stored
reference output in tests/golden/negsamp_fixture.npz.   gpu: the CUDA sampler through the C ABI against
the oracle's checker, a chi-square test against uniformity over the available nodes (same statistic
computed for the reference's stored output), determinism in the seed, error behaviour."""
import glob
import importlib.util
import os

import numpy as np
import pytest
import torch

from oracle import negative_sampling as ON

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FIX = np.load(os.path.join(ROOT, "tests", "golden", "negsamp_fixture.npz"))


def _compiled_reference():
    so = glob.glob(os.path.join(ROOT, "oracle", "_ref", "negative_sampling*.so"))
    if not so:
        return None
    spec = importlib.util.spec_from_file_location("negative_sampling", so[0])
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _chi2_z(drawn, n_nodes, bins=64):
    """z-score of the chi-square statistic of the drawn node ids binned into `bins` equal ranges
    (the unavailable set of a positive is ~1e-3 of the nodes: the available set is uniform to that order)."""
    counts = np.bincount((drawn.reshape(-1) * bins // n_nodes).astype(np.int64), minlength=bins).astype(np.float64)
    exp = drawn.size / bins
    chi2 = ((counts - exp) ** 2 / exp).sum()
    return (chi2 - (bins - 1)) / np.sqrt(2 * (bins - 1))


def test_reference_fixture_output_satisfies_the_checker():
    ei, pos = FIX["edge_index"], FIX["pos_edge_index"]
    drawn = ON.check(ei, pos, FIX["ref_neg"], 64)
    assert drawn.shape == (200, 64)
    assert abs(_chi2_z(drawn, 12797)) < 5


def test_oracle_port_matches_reference_behaviour():
    ei, pos = FIX["edge_index"], FIX["pos_edge_index"]
    neg = ON.generate_negative_samples(ei, pos, 64, np.random.default_rng(1))
    drawn = ON.check(ei, pos, neg, 64)
    assert abs(_chi2_z(drawn, 12797)) < 5
    with pytest.raises(ValueError):
        ON.generate_negative_samples(ei, pos, 0)
    assert np.asarray(ON.generate_negative_samples(ei, pos, 5)).shape == (2, 200 * 4)     # odd counts: 2 * (5 // 2)


def test_compiled_reference_agrees_with_the_checker():
    ref = _compiled_reference()
    if ref is None:
        pytest.skip("oracle/_ref not built (run oracle/build_ref.sh where /root/reference exists)")
    ei, pos = FIX["edge_index"], FIX["pos_edge_index"]
    neg = ref.generate_negative_samples(ei.tolist(), pos.tolist(), 8)
    ON.check(ei, pos, neg, 8)
    assert np.asarray(ref.generate_negative_samples(ei.tolist(), pos.tolist(), 5)).shape == (2, 200 * 4)
    with pytest.raises(ValueError):
        ref.generate_negative_samples(ei.tolist(), pos.tolist(), 0)


@pytest.mark.gpu
def test_cuda_sampler_parity():
    import b200mp
    ei, pos = torch.from_numpy(FIX["edge_index"]).cuda(), torch.from_numpy(FIX["pos_edge_index"]).cuda()
    neg = b200mp.generate_negative_samples(ei, pos, 64, seed=7)
    drawn = ON.check(FIX["edge_index"], FIX["pos_edge_index"], neg.cpu().numpy(), 64)
    z, z_ref = _chi2_z(drawn, 12797), _chi2_z(ON.check(FIX["edge_index"], FIX["pos_edge_index"], FIX["ref_neg"], 64), 12797)
    print(f"chi-square z: cuda {z:.2f}, compiled reference {z_ref:.2f}")
    assert abs(z) < 5
    assert torch.equal(neg, b200mp.generate_negative_samples(ei, pos, 64, seed=7))          # reproducible
    assert not torch.equal(neg, b200mp.generate_negative_samples(ei, pos, 64, seed=8))
    # list interface of the reference's pybind module, odd count, explicit node count
    out = b200mp.generate_negative_samples(FIX["edge_index"].tolist(), FIX["pos_edge_index"].tolist(), 5, seed=1)
    assert isinstance(out, list) and np.asarray(out).shape == (2, 800)
    ON.check(FIX["edge_index"], FIX["pos_edge_index"], out, 5)
    with pytest.raises(ValueError):
        b200mp.generate_negative_samples(ei, pos, 0)
    assert b200mp.generate_negative_samples(ei, pos[:, :0], 64).shape == (2, 0)


@pytest.mark.gpu
def test_cuda_sampler_dense_neighbourhoods():
    """hub positives: most nodes are neighbours, the few available ones must all be hit and nothing else"""
    import b200mp
    n = 40
    src = torch.zeros(n - 6, dtype=torch.long)
    dst = torch.arange(1, n - 5)
    extra = torch.tensor([[n - 5, n - 4, n - 3, n - 2, n - 1], [n - 4, n - 3, n - 2, n - 1, n - 5]])   # nodes not adjacent to 0
    ei = torch.cat([torch.stack([src, dst]), extra], 1)
    pos = torch.tensor([[0], [1]])
    neg = b200mp.generate_negative_samples(ei.cuda(), pos.cuda(), 512, seed=3).cpu().numpy()
    drawn = ON.check(ei.numpy(), pos.numpy(), neg, 512)
    assert set(drawn.reshape(-1).tolist()) == {n - 5, n - 4, n - 3, n - 2, n - 1}
    counts = np.bincount(drawn.reshape(-1) - (n - 5), minlength=5)
    assert counts.min() > 0.6 * 512 / 5
