"""-m gpu: the column-encoder kernels (csrc/col_encoder.cu) against the fp64 oracle
(oracle/stype_encoders.py; parity unpinned -- see its header): tokens within 1e-5 relative, parameter gradients
within 1e-5 relative of their fp64 values (fp32 reassociation of the sums over rows)."""
import pytest
import torch

from oracle import stype_encoders as O
from tests.enc_util import make_encoder, make_frame, oracle_blocks
from tests.util import rel_err

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("rows,C,ncat,cls", [(1000, 64, (15, 7), True), (70001, 64, (15, 7), False), (3000, 128, (300, 4), True),
                                              (513, 16, (3, 2), True), (1, 64, (5, 5), False)])
def test_tokens_and_parameter_gradients(rows, C, ncat, cls):
    tf, stats, names = make_frame(rows, seed=rows, ncat=ncat)
    enc = make_encoder(C, stats, names).cuda()
    tfd = tf.to("cuda")
    g = torch.Generator().manual_seed(7)
    cls_t = torch.randn(C, generator=g).cuda().requires_grad_(True) if cls else None
    go = torch.randn(rows, 5 + cls, C, generator=g)
    x, _ = enc(tfd, cls=cls_t)
    x.backward(go.cuda())
    blocks, P = oracle_blocks(enc, tf)
    ref = O.stype_wise_encode(blocks, None if cls_t is None else cls_t.detach().cpu().double())
    ref.backward(go.double())
    assert rel_err(x, ref) < 1e-5
    e = enc.encoder_dict
    pairs = {"num_w": (e["numerical"].weight.grad, P["num_w"].grad), "num_b": (e["numerical"].bias.grad, P["num_b"].grad),
             "ts_w": (e["timestamp"].weight.grad, P["ts_w"].grad), "ts_b": (e["timestamp"].bias.grad, P["ts_b"].grad)}
    for j, m in enumerate(e["categorical"].embs):
        assert float(m.weight.grad[0].abs().max()) == 0                       # padding row
        pairs[f"tab{j}"] = (m.weight.grad[1:], P["tabs"][j].grad[1:])
    if cls:
        pairs["cls"] = (cls_t.grad, go[:, 0].double().sum(0))
    for k, (a, b) in pairs.items():
        assert rel_err(a, b) < 1e-5, (k, rel_err(a, b))


def test_second_backward_and_two_timestamp_columns():
    tf, stats, names = make_frame(2000, seed=5, n_ts=2, n_num=3)
    enc = make_encoder(64, stats, names).cuda()
    tfd = tf.to("cuda")
    x, cols = enc(tfd)
    assert x.shape == (2000, 7, 64) and len(cols) == 7
    go = torch.randn_like(x)
    x.backward(go, retain_graph=True)
    g1 = enc.encoder_dict["timestamp"].weight.grad.clone()
    x.backward(go)
    assert rel_err(enc.encoder_dict["timestamp"].weight.grad, 2 * g1) < 1e-5
    blocks, P = oracle_blocks(enc, tf)
    assert rel_err(x, O.stype_wise_encode(blocks)) < 1e-5
