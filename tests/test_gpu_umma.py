"""-m gpu: the tcgen05 building blocks (csrc/umma.cuh) through b200mp_umma_selftest against fp64:
descriptor encodings, K-major / MN-major tile views, M=64 accumulator layout, A operand from
tensor memory, and the 3xTF32 split reaching fp32-level accuracy."""
import pytest
import torch

from b200mp import _lib as L
from tests.util import rel_err

pytestmark = pytest.mark.gpu


def run(mode, X, W):
    Y = torch.full((64 if mode in (2, 6) else 128, 64), float("nan"), device="cuda")
    L.check(L.lib().b200mp_umma_selftest(mode, L.ptr(X), L.ptr(W), L.ptr(Y), L.stream()), "umma_selftest")
    torch.cuda.synchronize()
    return Y


def truth(mode, X, W):
    X, W = X.double(), W.double()
    return X @ W if mode == 1 else X.t() @ W if mode in (2, 6) else X @ W.t()


@pytest.mark.parametrize("mode", [0, 3, 5, 6])
def test_umma_3xtf32(mode):
    g = torch.Generator(device="cuda").manual_seed(mode)
    X = torch.randn(128, 64, device="cuda", generator=g) * torch.logspace(-2, 2, 64, device="cuda")
    W = torch.randn(128 if mode in (2, 6) else 64, 64, device="cuda", generator=g)
    Y, T = run(mode, X, W), truth(mode, X, W)
    ref32 = truth(mode, X, W).float()
    err, err32 = rel_err(Y, T), rel_err(ref32, T)
    bias = float(((Y.double() - T) / T.abs().clamp(min=1e-3 * float(T.abs().max()))).mean())
    print(f"mode {mode}: 3xTF32 rel err {err:.2e} (fp32 torch matmul {err32:.2e}), mean signed err {bias:.2e}")
    assert err < 2e-6


def test_umma_single_pass_is_tf32():
    g = torch.Generator(device="cuda").manual_seed(9)
    X, W = torch.randn(128, 64, device="cuda", generator=g), torch.randn(64, 64, device="cuda", generator=g)
    err = rel_err(run(4, X, W), truth(4, X, W))
    print(f"single-pass tf32 rel err {err:.2e}")
    assert 1e-5 < err < 5e-3
