"""-m gpu: column attention: the attention-core kernels against
torch's scaled-dot-product attention in fp64, the whole layer (projections on the tcgen05 node kernels)
against nn.TransformerEncoderLayer itself in fp64, and the dropout mask's forward / backward consistency."""
import pytest
import torch
from torch import nn

from b200mp import functional as BF
from tests.util import rel_err

pytestmark = pytest.mark.gpu


def ref_core(q, k, v, nhead):
    R, S, C = q.shape
    hd = C // nhead
    sp = lambda t: t.view(R, S, nhead, hd).transpose(1, 2)
    a = torch.softmax(sp(q) @ sp(k).transpose(-1, -2) / hd ** 0.5, -1) @ sp(v)
    return a.transpose(1, 2).reshape(R, S, C)


@pytest.mark.parametrize("R,S,C,nhead", [(1000, 6, 64, 8), (333, 8, 64, 8), (257, 6, 128, 8), (64, 1, 64, 8), (100, 5, 32, 8), (90, 7, 16, 8), (50, 4, 8, 8)])
def test_core_fwd_bwd(R, S, C, nhead):
    g = torch.Generator(device="cuda").manual_seed(R + S)
    q, k, v, go = (torch.randn(R, S, C, device="cuda", generator=g) * 1.5 for _ in range(4))
    q, k, v = (t.requires_grad_(True) for t in (q, k, v))
    out = BF.col_attention(q, k, v, nhead)
    out.backward(go)
    qd, kd, vd = (t.detach().double().requires_grad_(True) for t in (q, k, v))
    od = ref_core(qd, kd, vd, nhead)
    od.backward(go.double())
    for nm, a, b in (("out", out, od), ("gq", q.grad, qd.grad), ("gk", k.grad, kd.grad), ("gv", v.grad, vd.grad)):
        err = rel_err(a, b)
        assert err < 1e-5, (nm, err)


def test_layer_matches_torch_encoder_layer():
    torch.manual_seed(0)
    R, S, C = 2000, 6, 64                       # R * S >= 4096: projections on the node kernels
    layer = nn.TransformerEncoderLayer(d_model=C, nhead=8, dim_feedforward=C, dropout=0.0, activation="relu", batch_first=True).cuda()
    x = torch.randn(R, S, C, device="cuda", requires_grad=True)
    go = torch.randn(R, S, C, device="cuda")
    y = BF.transformer_encoder_layer(layer, x)
    y.backward(go)
    grads = {n: p.grad.clone() for n, p in layer.named_parameters()}
    gx = x.grad.clone()
    ref = nn.TransformerEncoderLayer(d_model=C, nhead=8, dim_feedforward=C, dropout=0.0, activation="relu", batch_first=True).cuda().double()
    ref.load_state_dict({k_: v_.double() for k_, v_ in layer.state_dict().items()})
    xd = x.detach().double().requires_grad_(True)
    yd = ref(xd)
    yd.backward(go.double())
    worst = {"y": rel_err(y, yd), "gx": rel_err(gx, xd.grad)}
    for n, p in ref.named_parameters():
        worst[n] = rel_err(grads[n], p.grad)
    print({k_: f"{v_:.1e}" for k_, v_ in worst.items()})
    assert max(worst.values()) < 1e-5


def test_dropout_mask_consistent_between_forward_and_backward():
    R, S, C, nhead, p, seed = 500, 6, 64, 8, 0.5, 1234
    g = torch.Generator(device="cuda").manual_seed(3)
    q, k, v, go = (torch.randn(R, S, C, device="cuda", generator=g) for _ in range(4))
    out = BF._col_attention_fwd(q, k, v, nhead, p, seed)
    assert torch.equal(out, BF._col_attention_fwd(q, k, v, nhead, p, seed))
    assert not torch.equal(out, BF._col_attention_fwd(q, k, v, nhead, p, seed + 1))
    gq, gk, gv = BF._col_attention_bwd(q, k, v, go, nhead, p, seed)
    # out is linear in v for a fixed mask: <go, out(v)> == <gv, v>
    lhs, rhs = float((go.double() * out.double()).sum()), float((gv.double() * v.double()).sum())
    assert abs(lhs - rhs) < 1e-6 * abs(lhs) + 1e-3
    # directional derivative in q and k (same mask): central differences in fp64 accumulations
    dq, dk = torch.randn_like(q), torch.randn_like(k)
    eps = 1e-2
    f = lambda qq, kk: float((go.double() * BF._col_attention_fwd(qq, kk, v, nhead, p, seed).double()).sum())
    num = (f(q + eps * dq, k + eps * dk) - f(q - eps * dq, k - eps * dk)) / (2 * eps)
    ana = float((gq.double() * dq.double()).sum() + (gk.double() * dk.double()).sum())
    assert abs(num - ana) < 2e-2 * max(abs(ana), 1.0), (num, ana)
    # keep rate ~ 1 - p: E[out] preserved
    base = BF._col_attention_fwd(q, k, v, nhead, 0.0, 0)
    assert abs(float(out.mean()) - float(base.mean())) < 0.02


@pytest.mark.parametrize("R,S,C,nhead,p", [(1000, 6, 64, 8, 0.0), (333, 8, 64, 8, 0.3), (257, 6, 128, 8, 0.0), (100_003, 6, 64, 8, 0.2), (101, 5, 12, 3, 0.0),
                                           (77, 3, 16, 1, 0.1)])
def test_forward_loop_kernel_equals_one_group_per_cta(R, S, C, nhead, p):
    """B200MP_ATT_LOOP=1 (persistent CTAs, the next group's slices in flight during the arithmetic) is the same arithmetic per
    (row, head): bit for bit, dropout mask included, also where the group count is not a multiple of the groups per warp"""
    import os
    g = torch.Generator(device="cuda").manual_seed(R)
    q, k, v = (torch.randn(R, S, C, device="cuda", generator=g) for _ in range(3))
    outs = []
    for flag in ("0", "1"):
        os.environ["B200MP_ATT_LOOP"] = flag
        outs.append(BF._col_attention_fwd(q, k, v, nhead, p, 1234567))
    os.environ.pop("B200MP_ATT_LOOP")
    assert torch.equal(outs[0], outs[1])
