"""Host -> device input pipeline: double-buffered, asynchronous copies of one step's inputs from pinned host memory on
a copy stream, so that the transfer of step i + 1 overlaps the kernels of step i (the reference moves each batch with
blocking ``.to(device)`` calls inside the step: /root/reference/src/utils.py:346-356, fused.py:64-70).

    feeder = DeviceFeeder(device)
    feeder.stage(next_host_tensors)          # enqueue the copy of the NEXT step's inputs (returns at once)
    x, ei, ea = feeder.take()                # this step's device tensors; the compute stream waits for their copy

The two device buffer sets are allocated once per shape and re-used (no allocator traffic per step).
"""
from __future__ import annotations

from typing import Optional, Sequence, Tuple

import torch


class DeviceFeeder:
    def __init__(self, device: torch.device):
        if torch.device(device).type != "cuda":
            raise ValueError("DeviceFeeder copies to a CUDA device")
        self.device = torch.device(device)
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self._slots = [None, None]           # per slot: (device tensors, copy-done event, consumer-done event)
        self._staged: Optional[int] = None
        self._next = 0
        self._taken: Optional[int] = None
        self.bytes_staged = 0

    def _buffers(self, slot: int, host: Sequence[torch.Tensor]):
        cur = self._slots[slot]
        if cur is not None and all(d.shape == h.shape and d.dtype == h.dtype for d, h in zip(cur[0], host)) \
                and len(cur[0]) == len(host):
            return cur
        bufs = tuple(torch.empty(h.shape, dtype=h.dtype, device=self.device) for h in host)
        cur = (bufs, torch.cuda.Event(), torch.cuda.Event())
        self._slots[slot] = cur
        return cur

    def stage(self, host: Sequence[torch.Tensor]) -> None:
        """start copying ``host`` (pinned tensors) into the free buffer set"""
        if self._staged is not None:
            raise RuntimeError("a staged batch is waiting: call take() first")
        for h in host:
            if h.is_cuda or not h.is_pinned():
                raise ValueError("stage() takes pinned host tensors (asynchronous copies need page-locked memory)")
        slot = self._next
        bufs, copied, consumed = self._buffers(slot, host)
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(consumed)          # the step that last read this set has finished
            for d, h in zip(bufs, host):
                d.copy_(h, non_blocking=True)
                self.bytes_staged += h.numel() * h.element_size()
            copied.record(self.copy_stream)
        self._staged = slot
        self._next ^= 1

    def take(self) -> Tuple[torch.Tensor, ...]:
        """device tensors of the staged batch; valid until the second stage() after this call.  Call ``release()``
        once the step's kernels are enqueued so that the set can be overwritten."""
        if self._staged is None:
            raise RuntimeError("nothing staged")
        slot, self._staged = self._staged, None
        bufs, copied, _ = self._slots[slot]
        torch.cuda.current_stream(self.device).wait_event(copied)
        self._taken = slot
        return bufs

    def release(self) -> None:
        """mark the taken set as consumed at this point of the current stream"""
        if self._taken is None:
            raise RuntimeError("nothing taken")
        self._slots[self._taken][2].record(torch.cuda.current_stream(self.device))

    def drain(self) -> None:
        """make the current stream wait for every copy issued so far"""
        torch.cuda.current_stream(self.device).wait_stream(self.copy_stream)
