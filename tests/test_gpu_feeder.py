"""-m gpu: the double-buffered host -> device feeder returns exactly the staged bytes, in order, while copies of
the next batch overlap work on the current one."""
import pytest
import torch

import b200mp

pytestmark = pytest.mark.gpu


def test_feeder_round_trip_and_errors():
    dev = torch.device("cuda", 0)
    f = b200mp.DeviceFeeder(dev)
    batches = [(torch.randn(100000, 64).pin_memory(), torch.randint(0, 1000, (2, 100000)).pin_memory()) for _ in range(5)]
    with pytest.raises(RuntimeError):
        f.take()
    with pytest.raises(ValueError):
        f.stage((torch.zeros(4),))                      # not pinned
    f.stage(batches[0])
    with pytest.raises(RuntimeError):
        f.stage(batches[1])                             # one staged batch at a time
    sums = []
    for i in range(5):
        x, ei = f.take()
        if i + 1 < 5:
            f.stage(batches[i + 1])
        y = x
        for _ in range(20):                             # keep the compute stream busy while the next copy runs
            y = y * 1.0001 + 0.0
        assert torch.equal(x.cpu(), batches[i][0]) and torch.equal(ei.cpu(), batches[i][1])
        sums.append(float(y.sum()))
        f.release()
    f.drain()
    assert f.bytes_staged == sum(a.numel() * 4 + b.numel() * 8 for a, b in batches)


def test_feeder_varying_shapes_keep_their_bytes():
    """sampled batches change shape every step (ibm_transactions_for_aml.py:61-112): the two arenas grow a few times and
    are then re-used as views; no batch may be overwritten while the step that reads it is still running"""
    dev = torch.device("cuda", 0)
    f = b200mp.DeviceFeeder(dev)
    g = torch.Generator().manual_seed(0)
    sizes = [int(s) for s in torch.randint(50_000, 400_000, (12,), generator=g)]
    batches = [(torch.randn(s, 64, generator=g).pin_memory(), torch.randint(0, 1000, (2, s), generator=g).pin_memory(),
                torch.randn(s // 7, 3, generator=g).pin_memory()) for s in sizes]
    f.stage(batches[0])
    outs = []
    for i, b in enumerate(batches):
        x, ei, y = f.take()
        if i + 1 < len(batches):
            f.stage(batches[i + 1])
        acc = x
        for _ in range(30):                              # long-running consumer of THIS batch while the next copy is in flight
            acc = acc * 1.0001 + 0.0
        outs.append((acc.sum(), x.double().sum(), ei.sum(), y.double().sum()))
        f.release()
    f.drain()
    torch.cuda.synchronize()
    for (acc, sx, sei, sy), (hx, hei, hy) in zip(outs, batches):
        assert float(sx) == pytest.approx(float(hx.double().sum()), rel=1e-12, abs=1e-9)
        assert int(sei) == int(hei.sum()) and float(sy) == pytest.approx(float(hy.double().sum()), rel=1e-12, abs=1e-9)
    assert f.reallocations <= 2 * 6                        # arenas grow geometrically, they are not rebuilt per batch
