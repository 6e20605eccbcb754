"""bf16 STORAGE of the edge tensors (BASELINE.json config 3; north_star: 2e-2 for bf16): ``model.set_edge_storage(torch.bfloat16)``.
CPU: host logic with the emulated kernels (same rounding points: every [E,H] tensor rounded to bf16 where it is stored).
-m gpu: the bf16 kernels against that emulation (tight: same rounding points, fp32 arithmetic) and against the fp32 oracle
(2e-2 on the forward values).  Gradients of a piecewise-smooth network under bf16 storage: rounding e to 8 mantissa bits moves
~0.3 % of the relu pre-activations across their kink, each flip changes a gradient element by O(1), so the relative L2 distance to
the fp32 gradient is ~5e-2 for ANY bf16 implementation (measured with the plain-PyTorch emulation); the test bounds it at 1e-1 and
asks for a cosine similarity above 0.995."""
import pytest
import torch

import b200mp
from oracle import models as OM
from tests.emulator import emulated_kernels

H = 64


def _case(kind, n, e, seed):
    g = torch.Generator().manual_seed(seed)
    ei = torch.randint(0, n, (2, e), generator=g)
    ei[1, : e // 10] = 3                                     # a hub
    deg = torch.bincount(torch.bincount(ei[1], minlength=n))
    torch.manual_seed(seed)
    m = (b200mp.PNAS(num_features=H, num_gnn_layers=2, n_hidden=H, edge_dim=H, deg=deg) if kind == "pna"
         else b200mp.GINe(num_features=H, num_gnn_layers=2, n_hidden=H, edge_updates=True, edge_dim=H)).train()
    x, ea = torch.randn(n, H, generator=g), torch.randn(e, H, generator=g)
    wx, we = torch.randn(n, H, generator=g), torch.randn(e, H, generator=g)
    return m, ei, x, ea, wx, we


def _run(m, ei, x, ea, wx, we, dtype, device):
    m.set_edge_storage(dtype)
    m.zero_grad(set_to_none=True)
    xi, ei2 = x.detach().clone().to(device).requires_grad_(True), ea.detach().clone().to(device).requires_grad_(True)
    ox, oe = m.message_passing(xi, ei.to(device), ei2)
    assert oe.dtype == dtype and ox.dtype == torch.float32
    ((ox * wx.to(device)).sum() + (oe.float() * we.to(device)).sum()).backward()
    gw = torch.cat([p.grad.reshape(-1) for p in m.parameters() if p.grad is not None])
    return [t.detach().float().cpu() for t in (ox, oe, xi.grad, ei2.grad, gw)]


def _l2(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm())


def _cos(a, b):
    return float((a.double() * b.double()).sum() / (a.double().norm() * b.double().norm()))


@pytest.mark.parametrize("kind", ["pna", "gine"])
def test_bf16_storage_host_logic(kind):
    m, ei, x, ea, wx, we = _case(kind, 500, 6000, 3)
    with emulated_kernels():
        a = _run(m, ei, x, ea, wx, we, torch.float32, "cpu")
        b = _run(m, ei, x, ea, wx, we, torch.bfloat16, "cpu")
    assert _l2(b[0], a[0]) < 2e-2 and _l2(b[1], a[1]) < 2e-2
    for u, v in zip(b[2:], a[2:]):
        assert _l2(u, v) < 1e-1 and _cos(u, v) > 0.995


@pytest.mark.gpu
@pytest.mark.parametrize("kind,n,e", [("pna", 500, 6000), ("gine", 500, 6000), ("pna", 20_000, 300_001), ("gine", 20_000, 300_001)])
def test_bf16_kernels(kind, n, e):
    m, ei, x, ea, wx, we = _case(kind, n, e, 5)
    with emulated_kernels():
        emu = _run(m, ei, x, ea, wx, we, torch.bfloat16, "cpu")
    # fp32 oracle (CPU) for the forward values
    p = {k: v.detach().clone() for k, v in m.state_dict().items()}
    ox32, oe32 = OM.mp_stack(p, kind, x, ei, ea, 2, True, False, True)
    m = m.cuda()
    got = _run(m, ei, x, ea, wx, we, torch.bfloat16, "cuda")
    names = ("x'", "e'", "dL/dx", "dL/de", "dL/dW")
    print(f"bf16 {kind} N={n} E={e}: vs emulation " + ", ".join(f"{nm} {_l2(u, v):.1e}" for nm, u, v in zip(names, got, emu))
          + f" | vs fp32 oracle x' {_l2(got[0], ox32):.1e} e' {_l2(got[1], oe32):.1e}")
    # same rounding points, same arithmetic: forward values agree to a few bf16 ulps of single elements
    assert _l2(got[0], emu[0]) < 2e-3 and _l2(got[1], emu[1]) < 2e-3
    for u, v in zip(got[2:], emu[2:]):                        # gradients: kink decisions of values within fp32 round-off differ
        assert _l2(u, v) < 3e-2 and _cos(u, v) > 0.999
    assert _l2(got[0], ox32) < 2e-2 and _l2(got[1], oe32) < 2e-2
    rel = lambda a, b: float((a.double() - b.double()).abs().max() / b.double().abs().max())
    assert rel(got[0], ox32) < 2e-2 and rel(got[1], oe32) < 2e-2
