"""GPU debugging aid (round 2): where do the well-conditioned gradients of the H = 64 / N >= 4096 composites lose accuracy?"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200mp
from b200mp import functional as BF
from b200mp import graph as G
from tests.test_gpu_ops import hub_graph
from tests.util import rel_err

dev = "cuda"


def l2(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).norm() / b.norm())


def linear_group_case(n_w, N=20000):
    g = torch.Generator(device=dev).manual_seed(n_w)
    x = torch.randn(N, 64, device=dev, generator=g, requires_grad=True)
    Ws = [(torch.randn(64, 64, device=dev, generator=g) / 8).requires_grad_(True) for _ in range(n_w)]
    gy = [torch.randn(N, 64, device=dev, generator=g) for _ in range(n_w)]
    ys = BF.linear_group(x, Ws, [None] * n_w)
    torch.autograd.backward(ys, gy)
    xd = x.detach().double().requires_grad_(True)
    Wd = [w.detach().double().requires_grad_(True) for w in Ws]
    torch.autograd.backward([xd @ w.t() for w in Wd], [t.double() for t in gy])
    print(f"linear_group n={n_w}: y {l2(ys[0], xd @ Wd[0].t()):.2e} gx {l2(x.grad, xd.grad):.2e} gW {max(l2(a.grad, b.grad) for a, b in zip(Ws, Wd)):.2e}")


def pair_gather_case():
    n, e, W = 20000, 200000, 64
    ei = hub_graph(n, e, 3, 5000).to(dev)
    g = torch.Generator(device=dev).manual_seed(1)
    Rs, Rd, T = (torch.randn(r, W, device=dev, generator=g, requires_grad=True) for r in (n, n, e))
    w = torch.randn(e, W, device=dev, generator=g)
    out = BF.pair_gather_act(Rs, Rd, T, ei, True)
    (out * w).sum().backward()
    Rs64, Rd64, T64 = (t.detach().double().requires_grad_(True) for t in (Rs, Rd, T))
    o64 = torch.relu(Rs64[ei[0]] + Rd64[ei[1]] + T64)
    (o64 * w.double()).sum().backward()
    print(f"pair_gather_act: out {l2(out, o64):.2e} gRs {l2(Rs.grad, Rs64.grad):.2e} (max-norm {rel_err(Rs.grad, Rs64.grad):.2e}) "
          f"gRd {l2(Rd.grad, Rd64.grad):.2e} (max-norm {rel_err(Rd.grad, Rd64.grad):.2e}) gT {l2(T.grad, T64.grad):.2e}")


def tab_case(model, env):
    import tests.test_gpu_parity_scale as T
    for k, v in env.items():
        os.environ[k] = v
    G.clear_cache()
    T.REPORT = "/dev/null"
    try:
        T.test_tab_models_h64_vs_oracle(model)
        print(f"{model} {env}: PASS")
    except AssertionError as ex:
        print(f"{model} {env}: FAIL {str(ex)[:230]}")
    for k in env:
        os.environ.pop(k)


def wgrad_chain_case():
    """error of the edge update's W2 gradient (g^T h over E edges) against fp64 as a function of the drain interval"""
    E, n = 1_000_000, 65114
    g = torch.Generator(device=dev).manual_seed(3)
    ei = torch.stack([torch.randint(0, n, (E,), device=dev, generator=g), torch.randint(0, n, (E,), device=dev, generator=g)])
    st = G.edge_structure(ei, n)
    gout, e = torch.randn(E, 64, device=dev, generator=g), torch.randn(E, 64, device=dev, generator=g)
    h = torch.relu(torch.randn(E, 64, device=dev, generator=g))
    W1, W2 = torch.randn(64, 64, device=dev, generator=g) / 8, torch.randn(64, 64, device=dev, generator=g) / 8
    ref2 = 0.5 * gout.double().t() @ h.double()
    for fl in ("0", "8192", "2048", "1024", "512", "256"):
        os.environ["B200MP_WGRAD_FLUSH"] = fl
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        out = BF._edge_update_bwd(st, gout, h, e, W1, W2, 1.0, 0.5, True, True)
        t0.record()
        for _ in range(5):
            out = BF._edge_update_bwd(st, gout, h, e, W1, W2, 1.0, 0.5, True, True)
        t1.record()
        torch.cuda.synchronize()
        print(f"wgrad drain every {fl:>5s} edges: gW2 relL2 vs fp64 {l2(out[5], ref2):.2e}, {t0.elapsed_time(t1) / 5:.3f} ms per launch (E = 1 M)")
    os.environ.pop("B200MP_WGRAD_FLUSH")


if __name__ == "__main__":
    wgrad_chain_case()
    for n_w in (1, 2, 3):
        linear_group_case(n_w)
    pair_gather_case()
    for env in ({}, {"B200MP_FUSED_REDUCE": "0"}, {"B200MP_UPD_PIPE": "0"}, {"B200MP_FUSED_REDUCE": "0", "B200MP_UPD_PIPE": "0"},
                {"B200MP_SIMT": "1"}):
        tab_case("interleaved", env)
    tab_case("tabgnn", {"B200MP_FUSED_REDUCE": "0", "B200MP_UPD_PIPE": "0"})
