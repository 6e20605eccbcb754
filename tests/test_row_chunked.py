"""CPU (emulated kernels): functional.row_chunked -- chunk-by-chunk evaluation of the row-local column-attention layer with
recompute in backward -- gives the values and gradients of the plain evaluation, including with dropout active (the masks of
the recomputation are the masks of the forward run: seeds come from torch's generator, whose state is replayed per chunk)."""
import torch

from b200mp import functional as BF
from b200mp.nn.tab import FTTransformerLayer
from tests.emulator import emulated_kernels
from tests.util import rel_err


def _run(layer, x, w, chunk, seed):
    layer.chunk_rows = chunk
    layer.zero_grad(set_to_none=True)
    xi = x.clone().requires_grad_(True)
    torch.manual_seed(seed)
    y = layer(xi)
    (y * w).sum().backward()
    return y.detach(), xi.grad, [p.grad.clone() for p in layer.parameters()]


def test_chunked_equals_plain_and_replays_dropout():
    torch.manual_seed(0)
    R, S, C = 103, 4, 16
    x, w = torch.randn(R, S, C), torch.randn(R, S, C)
    with emulated_kernels():
        layer = FTTransformerLayer(C, nhead=4, dropout=0.0).train()
        a = _run(layer, x, w, 0, 1)
        b = _run(layer, x, w, 25, 1)
        assert torch.equal(a[0], b[0]) and rel_err(b[1], a[1]) < 1e-6
        assert all(rel_err(q, p) < 1e-5 for p, q in zip(a[2], b[2]))

        # dropout on: reference = the same chunk sequence evaluated WITH the graph (one generator draw per mask, same order)
        layer = FTTransformerLayer(C, nhead=4, dropout=0.3).train()
        c = _run(layer, x, w, 25, 7)
        layer.zero_grad(set_to_none=True)
        layer.chunk_rows = 0
        xi = x.clone().requires_grad_(True)
        torch.manual_seed(7)
        y = torch.cat([layer(xi[r:r + 25]) for r in range(0, R, 25)])
        (y * w).sum().backward()
        assert torch.equal(c[0], y.detach())
        assert rel_err(c[1], xi.grad) < 1e-6
        assert all(rel_err(q, p.grad) < 1e-5 for p, q in zip(layer.parameters(), c[2]))
        d = _run(layer, x, w, 25, 8)
        assert not torch.equal(c[0], d[0])                  # another seed, other masks


def test_manual_seed_reproduces_masks():
    with emulated_kernels():
        x = torch.randn(64, 32)
        torch.manual_seed(3)
        a = BF.relu_dropout(x, 0.5)
        torch.manual_seed(3)
        b = BF.relu_dropout(x, 0.5)
        c = BF.relu_dropout(x, 0.5)
        assert torch.equal(a, b) and not torch.equal(b, c)
