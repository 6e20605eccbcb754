"""-m gpu: the tcgen05 node-level linear kernels (csrc/node_gemm_umma.cu) against fp64: forward,
data gradient (transposed weight) and weight gradient, at ragged row counts and every supported
width; tolerance = fp32 round-off level (3xTF32), far below the 1e-5 the path needs."""
import pytest
import torch

from b200mp import functional as BF
from tests.util import rel_err

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("N,K,NO", [(4096, 256, 192), (5000, 192, 256), (4097, 128, 128), (9999, 256, 256), (5003, 384, 64), (4999, 512, 384), (4100, 384, 512)])
def test_node_linear(N, K, NO):
    g = torch.Generator(device="cuda").manual_seed(N + K)
    X = torch.randn(N, K, device="cuda", generator=g) * torch.logspace(-2, 2, K, device="cuda")
    W = torch.randn(NO, K, device="cuda", generator=g)
    b = torch.randn(NO, device="cuda", generator=g)
    Y = BF._node_linear(X, W, b)
    T = X.double() @ W.double().t() + b.double()
    err, err32 = rel_err(Y, T), rel_err(torch.nn.functional.linear(X, W, b), T)
    print(f"node_linear N={N} K={K} NO={NO}: rel err {err:.2e} (torch fp32 {err32:.2e})")
    assert err < 2e-6


@pytest.mark.parametrize("N,K,NO", [(4096, 256, 192), (5001, 128, 64), (7777, 256, 256), (4999, 512, 384), (4100, 384, 512)])
def test_node_wgrad(N, K, NO):
    g = torch.Generator(device="cuda").manual_seed(N + NO)
    X = torch.randn(N, K, device="cuda", generator=g)
    Y = torch.randn(N, NO, device="cuda", generator=g) * torch.logspace(-1, 1, NO, device="cuda")
    G = BF._node_wgrad(X, Y)
    T = X.double().t() @ Y.double()
    err, err32 = rel_err(G, T), rel_err(X.t() @ Y, T)
    print(f"node_wgrad N={N} K={K} NO={NO}: rel err {err:.2e} (torch fp32 {err32:.2e})")
    assert err < 2e-6


@pytest.mark.parametrize("N,K,NO", [(6000, 256, 192), (5000, 512, 384), (4500, 384, 64), (4500, 128, 128)])
def test_linear3x_autograd_matches_fp64(N, K, NO):
    """(5000, 512, 384): PNAConv post_nn at H = 128 -- more than 256 outputs, column blocks of one output (b200mp_node_linear_ld);
    (4500, 384, 64): the [E, K C] -> [E, H] edge embedding; (4500, 128, 128): the H = 128 node linears"""
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn(N, K, device="cuda", generator=g, requires_grad=True)
    W = (torch.randn(NO, K, device="cuda", generator=g) / 16).requires_grad_(True)
    b = torch.randn(NO, device="cuda", generator=g, requires_grad=True)
    gy = torch.randn(N, NO, device="cuda", generator=g)
    y = BF.linear3x(x, W, b)
    y.backward(gy)
    xd, Wd, bd = (t.detach().double().requires_grad_(True) for t in (x, W, b))
    yd = torch.nn.functional.linear(xd, Wd, bd)
    yd.backward(gy.double())
    for name, a, t in (("y", y, yd), ("gx", x.grad, xd.grad), ("gW", W.grad, Wd.grad), ("gb", b.grad, bd.grad)):
        err = rel_err(a, t)
        print(f"linear3x {name}: rel err {err:.2e}")
        assert err < 1e-5


@pytest.mark.parametrize("N", [4096, 7001])
def test_linear_group_matches_fp64(N):
    g = torch.Generator(device="cuda").manual_seed(N)
    x = torch.randn(N, 64, device="cuda", generator=g, requires_grad=True)
    Wbig = (torch.randn(64, 192, device="cuda", generator=g) / 8).requires_grad_(True)     # sliced like pre_nn's weight
    b2 = torch.randn(64, device="cuda", generator=g, requires_grad=True)
    Ws = (Wbig[:, :64], Wbig[:, 64:128], Wbig[:, 128:])
    gy = [torch.randn(N, 64, device="cuda", generator=g) for _ in range(3)]
    ys = BF.linear_group(x, Ws, (None, None, b2))
    torch.autograd.backward(ys, gy)
    xd, Wd, bd = (t.detach().double().requires_grad_(True) for t in (x, Wbig, b2))
    yd = [xd @ Wd[:, :64].t(), xd @ Wd[:, 64:128].t(), xd @ Wd[:, 128:].t() + bd]
    torch.autograd.backward(yd, [t.double() for t in gy])
    for name, a, t in (("y2", ys[2], yd[2]), ("gx", x.grad, xd.grad), ("gW", Wbig.grad, Wd.grad), ("gb", b2.grad, bd.grad)):
        err = rel_err(a, t)
        print(f"linear_group N={N} {name}: rel err {err:.2e}")
        assert err < 1e-5


@pytest.mark.parametrize("N,n", [(4096, 3), (7001, 2), (5000, 1)])
def test_linear_cat_matches_fp64(N, n):
    """y = [x_0 | x_1 | x_2] W^T + b without the concatenation (reverse-message-passing `lin`): values and every gradient vs fp64"""
    g = torch.Generator(device="cuda").manual_seed(N + n)
    xs = [torch.randn(N, 64, device="cuda", generator=g, requires_grad=True) for _ in range(n)]
    W = (torch.randn(64, 64 * n, device="cuda", generator=g) / 8).requires_grad_(True)
    b = torch.randn(64, device="cuda", generator=g, requires_grad=True)
    gy = torch.randn(N, 64, device="cuda", generator=g)
    y = BF.linear_cat(xs, W, b)
    y.backward(gy)
    xd = [x.detach().double().requires_grad_(True) for x in xs]
    Wd, bd = W.detach().double().requires_grad_(True), b.detach().double().requires_grad_(True)
    yd = torch.cat(xd, 1) @ Wd.t() + bd
    yd.backward(gy.double())
    pairs = [("y", y, yd), ("gW", W.grad, Wd.grad), ("gb", b.grad, bd.grad)] + [(f"gx{i}", a.grad, c.grad) for i, (a, c) in enumerate(zip(xs, xd))]
    for name, a, t in pairs:
        err = rel_err(a, t)
        print(f"linear_cat N={N} n={n} {name}: rel err {err:.2e}")
        assert err < 1e-5


@pytest.mark.parametrize("N,n", [(4096, 3), (7001, 2), (130, 1), (300_001, 3)])
def test_node_linear64_colsum(N, n):
    """data gradient of a group of 64 -> 64 linears and their bias gradients (column sums of the incoming gradients) out of one pass"""
    g = torch.Generator(device="cuda").manual_seed(N + n)
    xs = [torch.randn(N, 64, device="cuda", generator=g) * (1 + i) + 0.25 * i for i in range(n)]
    W = torch.randn(64, 64 * n, device="cuda", generator=g) / 8
    y, cs = BF._node_linear64_colsum(xs, W)
    X = torch.cat(xs, 1).double()
    assert rel_err(y, X @ W.double().t()) < 2e-6
    assert rel_err(cs, X.sum(0)) < 2e-6
    y2, cs2 = BF._node_linear64_colsum(xs, W)            # fixed summation order: bit-identical from run to run
    assert torch.equal(y, y2) and torch.equal(cs, cs2)
    assert torch.equal(y, BF._node_linear64(xs, W, None, False))


@pytest.mark.parametrize("N", [130, 4096, 300_001])
def test_node_pair_wgrad_single_gradient(N):
    """b200mp_node_pair_wgrad with G1 == NULL: g^T x with g read once, the second half of the partials zero"""
    g = torch.Generator(device="cuda").manual_seed(N)
    g0 = torch.randn(N, 64, device="cuda", generator=g)
    x = torch.randn(N, 64, device="cuda", generator=g)
    a, z = BF._node_pair_wgrad(g0, None, x)
    assert rel_err(a, g0.double().t() @ x.double()) < (2e-6 if N < 100_000 else 5e-6)      # a sum over 300 k rows: 3.4e-6 measured
    assert not bool(z.any())
    b, _ = BF._node_pair_wgrad(g0, g0, x)
    assert torch.equal(a, b)
