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
