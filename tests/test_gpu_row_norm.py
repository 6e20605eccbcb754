"""-m gpu: residual + dropout + LayerNorm kernels (csrc/row_norm.cu) against torch's LayerNorm in fp64, and the
dropout mask's forward / backward consistency."""
import pytest
import torch
from torch import nn

from b200mp import functional as BF
from tests.util import rel_err

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("rows,C,with_sub", [(1000, 64, True), (70001, 64, False), (513, 16, True), (300, 32, True), (999, 128, True), (1, 64, False)])
def test_against_torch_layernorm(rows, C, with_sub):
    g = torch.Generator(device="cuda").manual_seed(rows + C)
    x = (torch.randn(rows, C, device="cuda", generator=g) * 3 + 1).requires_grad_(True)
    sub = torch.randn(rows, C, device="cuda", generator=g).requires_grad_(True) if with_sub else None
    norm = nn.LayerNorm(C).cuda()
    with torch.no_grad():
        norm.weight.normal_(1.0, 0.3, generator=g)
        norm.bias.normal_(0.0, 0.3, generator=g)
    go = torch.randn(rows, C, device="cuda", generator=g)
    y = BF.add_dropout_layernorm(x, sub, 0.0, norm)
    y.backward(go)
    ref = nn.LayerNorm(C).cuda().double()
    ref.load_state_dict({k: v.double() for k, v in norm.state_dict().items()})
    xd = x.detach().double().requires_grad_(True)
    sd = sub.detach().double().requires_grad_(True) if with_sub else None
    yd = ref(xd if sd is None else xd + sd)
    yd.backward(go.double())
    errs = {"y": rel_err(y, yd), "gx": rel_err(x.grad, xd.grad), "gw": rel_err(norm.weight.grad, ref.weight.grad),
            "gb": rel_err(norm.bias.grad, ref.bias.grad)}
    if with_sub:
        errs["gsub"] = rel_err(sub.grad, sd.grad)
    assert max(errs.values()) < 1e-5, errs


def test_dropout_mask_is_the_same_in_forward_and_backward():
    rows, C, p, seed = 20000, 64, 0.5, 99
    g = torch.Generator(device="cuda").manual_seed(1)
    x, sub, go = (torch.randn(rows, C, device="cuda", generator=g) for _ in range(3))
    gamma, beta = torch.rand(C, device="cuda", generator=g) + 0.5, torch.randn(C, device="cuda", generator=g)
    y, stats = BF._row_norm_fwd(x, sub, p, seed, gamma, beta, 1e-5)
    gx, gsub, gg, gb = BF._row_norm_bwd(go, x, sub, stats, p, seed, gamma)
    ok = gx.abs() > 1e-6
    ratio = (gsub / gx)[ok]
    assert bool(((ratio.abs() < 1e-6) | ((ratio - 2.0).abs() < 1e-4)).all())
    keep = float((ratio > 1.0).float().mean())
    assert abs(keep - (1 - p)) < 0.01
    mask = torch.where(ok, (gsub / gx).round(), torch.full_like(gx, 2.0))      # undetermined entries: gx ~ 0, rare
    z = (x + sub * mask).double()
    ref = torch.nn.functional.layer_norm(z, (C,), gamma.double(), beta.double(), 1e-5)
    bad = (~ok).any(dim=1)
    assert rel_err(y[~bad], ref[~bad]) < 1e-5
    y2, _ = BF._row_norm_fwd(x, sub, p, seed + 1, gamma, beta, 1e-5)
    assert not torch.equal(y, y2)


def test_relu_dropout():
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn(5000, 64, device="cuda", generator=g, requires_grad=True)
    go = torch.randn(5000, 64, device="cuda", generator=g)
    y = BF.relu_dropout(x, 0.0)
    y.backward(go)
    assert torch.equal(y, torch.relu(x.detach())) and torch.equal(x.grad, torch.where(x.detach() > 0, go, torch.zeros_like(go)))
    y5 = BF._relu_dropout(x.detach(), None, 0.5, 77)
    r = torch.relu(x.detach())
    assert bool(((y5 == 0) | (y5 == 2 * r)).all())
    kept = float((y5[r > 0] > 0).float().mean())
    assert abs(kept - 0.5) < 0.01
    gx = BF._relu_dropout(y5, go, 0.5, 77)
    assert torch.equal(gx, torch.where(y5 > 0, 2 * go, torch.zeros_like(go)))
    assert not torch.equal(y5, BF._relu_dropout(x.detach(), None, 0.5, 78))
