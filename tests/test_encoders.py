"""CPU: the stype-wise encoder drop-ins -- host logic over the emulated kernels
(tests/emulator.py -> oracle/stype_encoders.py), and the oracle against an independent scalar restatement."""
import math

import pytest
import torch

from b200mp.nn import (EmbeddingEncoder, LinearEncoder, ProjectionEncoder, StatType, StypeWiseFeatureEncoder, TensorFrame,
                       stype)
from oracle import stype_encoders as O
from tests.emulator import emulated_kernels
from tests.enc_util import make_encoder, make_frame, oracle_blocks
from tests.util import rel_err


def test_oracle_against_scalar_formulas():
    tf, stats, names = make_frame(7, seed=3)
    enc = make_encoder(16, stats, names)
    blocks, P = oracle_blocks(enc, tf)
    x = O.stype_wise_encode(blocks)
    assert x.shape == (7, 5, 16)
    num, cat, ts = (tf.feat_dict[s] for s in (stype.numerical, stype.categorical, stype.timestamp))
    e = enc.encoder_dict
    for r in range(7):
        for j in range(2):
            z = (float(num[r, j]) - float(e["numerical"].mean[j].double())) / float(e["numerical"].std[j].double())
            ref = z * P["num_w"][j] + P["num_b"][j]
            assert rel_err(x[r, j], ref) < 1e-12
            assert torch.equal(x[r, 2 + j], P["tabs"][j][int(cat[r, j]) + 1])
        f = [float(v) for v in ts[r, 0]]
        encs = []
        yr = f[0] - 2019
        encs += [math.sin(yr * 10 ** (-q)) for q in range(4)] + [math.cos(yr * 10 ** (-q)) for q in range(4)]
        for v, mx in zip(f[1:], O.CYCLIC_MAX):
            encs += [math.sin(v / mx * (q + 1) * math.pi) for q in range(4)] + [math.cos(v / mx * (q + 1) * 2 * math.pi) for q in range(4)]
        ref = torch.tensor(encs, dtype=torch.float64) @ P["ts_w"][0].reshape(56, -1) + P["ts_b"][0]
        assert rel_err(x[r, 4], ref) < 1e-9


def test_encoder_dropin_on_emulated_kernels():
    tf, stats, names = make_frame(50, seed=1)
    enc = make_encoder(16, stats, names)
    keys = set(enc.state_dict().keys())
    assert {"encoder_dict.numerical.weight", "encoder_dict.numerical.bias", "encoder_dict.numerical.mean", "encoder_dict.numerical.std",
            "encoder_dict.categorical.embs.0.weight", "encoder_dict.categorical.embs.1.weight", "encoder_dict.timestamp.weight",
            "encoder_dict.timestamp.bias", "encoder_dict.timestamp.min_year"} <= keys
    cls = torch.randn(16, requires_grad=True)
    go = torch.randn(50, 6, 16)
    with emulated_kernels():
        x, cols = enc(tf)
        assert x.shape == (50, 5, 16) and cols == names[stype.numerical] + names[stype.categorical] + names[stype.timestamp]
        xc, _ = enc(tf, cls=cls)
        assert torch.equal(xc[:, 1:], x) and torch.equal(xc[:, 0], cls.detach().expand(50, 16))
        xc.backward(go)
        sub, _ = enc(tf[[3, 4, 10]])
        assert torch.equal(sub, x[[3, 4, 10]])
    blocks, P = oracle_blocks(enc, tf, torch.float32)
    ref = O.stype_wise_encode(blocks, cls.detach())
    ref.backward(go)
    assert rel_err(xc, ref) < 1e-6
    e = enc.encoder_dict
    assert rel_err(e["numerical"].weight.grad, P["num_w"].grad) < 1e-5 and rel_err(e["timestamp"].weight.grad, P["ts_w"].grad) < 1e-5
    assert rel_err(cls.grad, go[:, 0].sum(0)) < 1e-5
    g0 = e["categorical"].embs[0].weight.grad
    assert float(g0[0].abs().max()) == 0 and rel_err(g0[1:], P["tabs"][0].grad[1:]) < 1e-5


def test_tensor_frame_and_errors():
    tf, stats, names = make_frame(20, seed=2)
    assert tf.num_cols == 5 and tf.num_rows == 20 and len(tf[torch.arange(5)]) == 5 and tf[2:7].y.shape == (5,)
    assert tf.get_col_feat("cat1").shape == (20, 1)
    with pytest.raises(ValueError):
        TensorFrame({stype.numerical: torch.zeros(3, 2)}, {stype.numerical: ["a"]})
    with pytest.raises(ValueError):                                     # a stype without an encoder
        StypeWiseFeatureEncoder(8, stats, names, {stype.numerical: LinearEncoder()})
    with pytest.raises(ValueError):                                     # wrong encoder for the stype
        StypeWiseFeatureEncoder(8, stats, {stype.numerical: names[stype.numerical]}, {stype.numerical: EmbeddingEncoder()})
    with pytest.raises(KeyError):
        StypeWiseFeatureEncoder(8, {"amount0": {}, "amount1": {}}, {stype.numerical: names[stype.numerical]}, {stype.numerical: LinearEncoder()})
    node_tf = TensorFrame({stype.relation: torch.ones(4, 1)}, {stype.relation: ["node_attr"]})
    enc = StypeWiseFeatureEncoder(8, {"node_attr": {}}, node_tf.col_names_dict, {stype.relation: ProjectionEncoder()})
    with emulated_kernels():
        x, _ = enc(node_tf)
    assert rel_err(x[:, 0], (enc.encoder_dict["relation"].weight + enc.encoder_dict["relation"].bias).expand(4, 8)) < 1e-6
