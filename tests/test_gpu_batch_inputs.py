"""-m gpu: device batch construction (b200mp.batch.BatchBuilder: lp_inputs / mcm_inputs / graph_inputs)
against the numpy restatement of src/utils/batch_processing.py (oracle/batch_inputs.py) in the take-all regime: node features,
relabelled edge_index, gathered / replicated attribute rows bit for bit; the negative edges by the reference's validity rule."""
import numpy as np
import pytest
import torch

import b200mp
from b200mp.batch import BatchBuilder
from b200mp.nn import TensorFrame, stype
from oracle import batch_inputs as OB
from oracle import index_ops as OI

pytestmark = pytest.mark.gpu


def _table(e, seed):
    g = torch.Generator().manual_seed(seed)
    names = {stype.numerical: ["a", "b"], stype.categorical: ["c"], stype.timestamp: ["t"]}
    feats = {stype.numerical: torch.randn(e, 2, generator=g), stype.categorical: torch.randint(0, 9, (e, 1), generator=g),
             stype.timestamp: torch.randint(0, 60, (e, 1, 7), generator=g)}
    return TensorFrame(feats, names), {str(k): v.numpy() for k, v in feats.items()}


def _eq_frame(tf, ref):
    for k, v in tf.feat_dict.items():
        assert np.array_equal(v.cpu().numpy(), ref[str(k)]), k


@pytest.mark.parametrize("n,e,b,ego", [(3000, 9000, 50, False), (800, 5000, 33, True), (200, 260, 5, True)])
def test_lp_and_mcm_inputs_bit_exact(n, e, b, ego):
    g = torch.Generator().manual_seed(n + e)
    ei = torch.randint(0, n, (2, e), generator=g)
    tf, feats = _table(e, 3)
    perm, colptr, row = OI.csc_build(ei.numpy(), n)
    ids = torch.randperm(e, generator=g)[:b]
    seeds = torch.stack([ei[0, ids], ei[1, ids], ids], 1)
    bb = BatchBuilder(ei.cuda(), n, tf.to("cuda"), (100, 100))
    table = ei.t().contiguous().numpy()

    nf, eidx, ea, nei, nea, tei, tea = bb.lp_inputs(seeds.cuda(), num_neg_samples=64, ego=ego, seed=7)
    r_nf, r_eidx, r_ea, r_nei, r_nea, r_pos, r_tea = OB.lp_inputs(table, perm, colptr, row, feats, seeds.numpy(), 64, (100, 100), ego)
    assert np.array_equal(nf.cpu().numpy(), r_nf) and np.array_equal(eidx.cpu().numpy(), r_eidx)
    assert np.array_equal(nei.cpu().numpy(), r_nei) and np.array_equal(tei[:, :b].cpu().numpy(), r_pos)
    _eq_frame(ea, r_ea); _eq_frame(nea, r_nea); _eq_frame(tea, r_tea)
    assert tea.num_rows == b * 65 and tei.shape == (2, b * 65)
    assert OI.negative_samples_valid(r_eidx, r_pos, tei[:, b:].cpu().numpy(), 64)
    again = bb.lp_inputs(seeds.cuda(), num_neg_samples=64, ego=ego, seed=7)
    assert torch.equal(again[5], tei)                                   # same seed, same negatives

    m = bb.mcm_inputs(seeds.cuda(), ego=ego)
    r = OB.mcm_inputs(table, perm, colptr, row, feats, seeds.numpy(), (100, 100), ego)
    assert np.array_equal(m[0].cpu().numpy(), r[0]) and np.array_equal(m[1].cpu().numpy(), r[1]) and np.array_equal(m[3].cpu().numpy(), r[3])
    _eq_frame(m[2], r[2]); _eq_frame(m[4], r[4])

    y = torch.randint(0, 2, (b,))
    gi = bb.graph_inputs(seeds.cuda(), y.cuda(), ego=ego)
    assert np.array_equal(gi[0].cpu().numpy(), r[0]) and np.array_equal(gi[1].cpu().numpy(), r[1]) and torch.equal(gi[3].cpu(), y)
