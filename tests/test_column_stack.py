"""The column-attention stack as one autograd node (b200mp/column_stack.py) against the layer-by-layer composition of the same
kernels (functional.transformer_encoder_layer + add_dropout_layernorm + torch's adds), which the oracle-backed tests pin to the
reference: values and every gradient, with and without dropout (seeds pinned), chunked and unchunked.  CPU: through the emulated
kernels; -m gpu: the real ones."""
import pytest
import torch

import b200mp
from b200mp import column_stack as CS
from b200mp import functional as BF
from tests.util import rel_err


def _layers(L, dropout, seed):
    torch.manual_seed(seed)
    convs = [torch.nn.TransformerEncoderLayer(d_model=64, nhead=8, dim_feedforward=64, dropout=dropout, activation="relu", batch_first=True)
             for _ in range(L)]
    norms = [torch.nn.LayerNorm(64) for _ in range(L)]
    for m in convs + norms:
        for p in m.parameters():
            if p.dim() == 1:
                torch.nn.init.normal_(p, mean=1.0 if p is getattr(m, "weight", None) else 0.0, std=0.3)
    return convs, norms


def _reference(convs, norms, x, add_input):
    t = x
    for c, n in zip(convs, norms):
        t = (t + BF.add_dropout_layernorm(BF.transformer_encoder_layer(c, t), None, 0.0, n)) / 2
    return x + t if add_input else t


def _run(fn, convs, norms, x, w):
    for m in convs + norms:
        m.zero_grad(set_to_none=True)
    x = x.detach().clone().requires_grad_(True)
    y = fn(x)
    (y * w).sum().backward()
    grads = [p.grad.detach().clone() for m in convs + norms for p in m.parameters()]
    return y.detach(), x.grad.detach(), grads


def _compare(dev, R, S, L, dropout, chunk, add_input, monkeypatch):
    convs, norms = _layers(L, dropout, 3 + L)
    for m in convs + norms:
        m.to(dev).train()
    g = torch.Generator().manual_seed(R + S)
    x = torch.randn(R, S, 64, generator=g).to(dev)
    w = torch.randn(R, S, 64, generator=g).to(dev)
    monkeypatch.setattr(CS, "MIN_TOKEN_ROWS", 1)
    if dropout > 0:                 # every dropout site draws the same seed in both paths; one chunk, so the masks' row indices agree
        monkeypatch.setattr(BF, "_next_seed", lambda salt: 0x5DEECE66D)
    assert CS.supported(convs, norms, x) or dev == "cpu"
    got = _run(lambda t: CS.column_stack(convs, norms, t, chunk, add_input), convs, norms, x, w)
    ref = _run(lambda t: _reference(convs, norms, t, add_input), convs, norms, x, w)
    assert rel_err(got[0], ref[0]) < 1e-5
    assert rel_err(got[1], ref[1]) < 1e-5
    names = [f"{i}.{k}" for i, m in enumerate(convs + norms) for k, _ in m.named_parameters()]
    for nme, a, b in zip(names, got[2], ref[2]):
        assert a.shape == b.shape and rel_err(a, b) < 2e-5, (nme, rel_err(a, b))


@pytest.mark.parametrize("L,dropout,chunk,add_input", [(1, 0.0, 0, False), (2, 0.0, 0, True), (2, 0.0, 7, True), (3, 0.0, 5, True),
                                                      (2, 0.3, 0, True)])
def test_column_stack_equals_layerwise_emulated(L, dropout, chunk, add_input, monkeypatch):
    from tests.emulator import emulated_kernels
    with emulated_kernels():
        _compare("cpu", 19, 6, L, dropout, chunk, add_input, monkeypatch)


@pytest.mark.gpu
@pytest.mark.parametrize("R,S,L,dropout,chunk,add_input", [(700, 6, 1, 0.0, 0, False), (5000, 6, 2, 0.0, 0, True), (5000, 6, 2, 0.0, 1024, True),
                                                          (3000, 2, 3, 0.0, 1000, True), (4000, 6, 2, 0.3, 0, True)])
def test_column_stack_equals_layerwise_gpu(R, S, L, dropout, chunk, add_input, monkeypatch):
    _compare("cuda", R, S, L, dropout, chunk, add_input, monkeypatch)


@pytest.mark.gpu
def test_column_stack_dropout_recompute_is_consistent():
    """chunked, with dropout: seeds are drawn once per chunk in forward and reused by the recompute -- two runs from one torch seed agree
    bit for bit (the LayerNorm gamma / beta sums excepted: vector REDs) and everything is finite"""
    convs, norms = _layers(2, 0.2, 9)
    for m in convs + norms:
        m.to("cuda").train()
    g = torch.Generator(device="cuda").manual_seed(1)
    x = torch.randn(3000, 6, 64, device="cuda", generator=g)
    w = torch.randn(3000, 6, 64, device="cuda", generator=g)
    outs = []
    for _ in range(2):
        torch.manual_seed(77)
        outs.append(_run(lambda t: CS.column_stack(convs, norms, t, 1024, True), convs, norms, x, w))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
    assert all(rel_err(a, b) < 1e-5 for a, b in zip(outs[0][2], outs[1][2]))
    assert all(bool(torch.isfinite(t).all()) for t in [outs[0][0], outs[0][1], *outs[0][2]])
