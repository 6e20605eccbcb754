"""-m gpu: fused BatchNorm + ReLU + residual average (csrc/node_norm.cu) against torch's BatchNorm1d in fp64:
outputs, all gradients and the running statistics."""
import pytest
import torch
from torch import nn

from b200mp import functional as BF
from tests.util import rel_err

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("N,H", [(1000, 64), (100001, 64), (777, 16), (5000, 32), (3001, 128)])
def test_bn_relu_residual(N, H):
    g = torch.Generator(device="cuda").manual_seed(N + H)
    c = (torch.randn(N, H, device="cuda", generator=g) * 2 + 0.7).requires_grad_(True)
    x = torch.randn(N, H, device="cuda", generator=g).requires_grad_(True)
    go = torch.randn(N, H, device="cuda", generator=g)
    bn = nn.BatchNorm1d(H).cuda()
    with torch.no_grad():
        bn.weight.normal_(1.0, 0.3, generator=g)
        bn.bias.normal_(0.0, 0.3, generator=g)
    ref = nn.BatchNorm1d(H).cuda().double()
    ref.load_state_dict({k: (v.double() if v.is_floating_point() else v.clone()) for k, v in bn.state_dict().items()})
    for step in range(2):                                   # two steps: running statistics accumulate
        y = BF.bn_relu_residual(bn, c, x)
    y.backward(go)
    cd, xd = c.detach().double().requires_grad_(True), x.detach().double().requires_grad_(True)
    for step in range(2):
        yd = (xd + torch.relu(ref(cd))) / 2
    yd.backward(go.double())
    errs = {"y": rel_err(y, yd), "gc": rel_err(c.grad, cd.grad), "gx": rel_err(x.grad, xd.grad),
            "gw": rel_err(bn.weight.grad, ref.weight.grad), "gb": rel_err(bn.bias.grad, ref.bias.grad),
            "rm": rel_err(bn.running_mean, ref.running_mean), "rv": rel_err(bn.running_var, ref.running_var)}
    assert max(errs.values()) < 1e-5, errs
    assert int(bn.num_batches_tracked) == 2
    bn.eval()                                               # eval mode: running statistics, torch path
    ref.eval()
    assert rel_err(BF.bn_relu_residual(bn, c, x), (xd + torch.relu(ref(cd))) / 2) < 1e-5
