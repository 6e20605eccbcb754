"""Host -> device input pipeline: double-buffered, asynchronous copies of one step's inputs from pinned host memory on
a copy stream, so that the transfer of step i + 1 overlaps the kernels of step i (the reference moves each batch with
blocking ``.to(device)`` calls inside the step: /root/reference/src/utils.py:346-356, fused.py:64-70).

    feeder = DeviceFeeder(device)
    feeder.stage(next_host_tensors)          # enqueue the copy of the NEXT step's inputs (returns at once)
    x, ei, ea = feeder.take()                # this step's device tensors; the compute stream waits for their copy
    ...enqueue the step...
    feeder.release()                         # the set may be overwritten once the step's kernels are done

Each of the two device buffer sets is ONE byte arena sized for the largest batch seen so far; the tensors handed out are
views into it, so the sampled batches of the reference (a different shape every step) do not reallocate.  Stream safety:
arenas are allocated under the copy stream and marked as used by the compute stream (``record_stream``); when an arena
has to grow, the copy stream first waits for everything the compute stream has enqueued so far (the step that may still
read the old arena), and the old arena's block is only recycled by the allocator after both streams have passed it.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch

_ALIGN = 256


class _Slot:
    __slots__ = ("arena", "views", "copied", "consumed", "used")

    def __init__(self):
        self.arena: Optional[torch.Tensor] = None
        self.views: Tuple[torch.Tensor, ...] = ()
        self.copied = torch.cuda.Event()
        self.consumed = torch.cuda.Event()
        self.used = False                      # consumed has been recorded at least once


class DeviceFeeder:
    def __init__(self, device: torch.device, capacity_bytes: int = 0):
        if torch.device(device).type != "cuda":
            raise ValueError("DeviceFeeder copies to a CUDA device")
        self.device = torch.device(device)
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self._slots: List[_Slot] = [_Slot(), _Slot()]
        self._staged: Optional[int] = None
        self._next = 0
        self._taken: Optional[int] = None
        self._capacity = int(capacity_bytes)
        self.bytes_staged = 0
        self.reallocations = 0

    @staticmethod
    def _layout(host: Sequence[torch.Tensor]):
        offs, total = [], 0
        for h in host:
            offs.append(total)
            total += (h.numel() * h.element_size() + _ALIGN - 1) // _ALIGN * _ALIGN
        return offs, total

    def _views(self, slot: _Slot, host: Sequence[torch.Tensor]) -> Tuple[torch.Tensor, ...]:
        """views of the slot's arena shaped like ``host`` (growing the arena when the batch does not fit); runs with the
        copy stream current"""
        offs, total = self._layout(host)
        compute = torch.cuda.current_stream(self.device)
        if slot.arena is None or slot.arena.numel() < total:
            cap = max(total, self._capacity, int(slot.arena.numel() * 1.5) if slot.arena is not None else 0)
            # the step that last read the old arena may still be running on the compute stream: order the new arena's first
            # copy after everything enqueued there so far (its memory may be the block the allocator recycles)
            ev = torch.cuda.Event()
            ev.record(compute)
            with torch.cuda.stream(self.copy_stream):
                self.copy_stream.wait_event(ev)
                slot.arena = torch.empty(cap, dtype=torch.uint8, device=self.device)
            slot.arena.record_stream(compute)                 # read by kernels of the compute stream
            self.reallocations += 1
        out = []
        for h, o in zip(host, offs):
            nb = h.numel() * h.element_size()
            out.append(slot.arena[o:o + nb].view(h.dtype).view(h.shape))
        return tuple(out)

    def stage(self, host: Sequence[torch.Tensor]) -> None:
        """start copying ``host`` (pinned tensors) into the free buffer set"""
        if self._staged is not None:
            raise RuntimeError("a staged batch is waiting: call take() first")
        for h in host:
            if h.is_cuda or not h.is_pinned():
                raise ValueError("stage() takes pinned host tensors (asynchronous copies need page-locked memory)")
            if not h.is_contiguous():
                raise ValueError("stage() takes contiguous host tensors")
        slot = self._slots[self._next]
        slot.views = self._views(slot, host)
        with torch.cuda.stream(self.copy_stream):
            if slot.used:
                self.copy_stream.wait_event(slot.consumed)     # the step that last read this set has finished
            for d, h in zip(slot.views, host):
                d.copy_(h, non_blocking=True)
                self.bytes_staged += h.numel() * h.element_size()
            slot.copied.record(self.copy_stream)
        self._staged = self._next
        self._next ^= 1

    def take(self) -> Tuple[torch.Tensor, ...]:
        """device tensors of the staged batch; valid until the second stage() after this call.  Call ``release()``
        once the step's kernels are enqueued so that the set can be overwritten."""
        if self._staged is None:
            raise RuntimeError("nothing staged")
        idx, self._staged = self._staged, None
        slot = self._slots[idx]
        torch.cuda.current_stream(self.device).wait_event(slot.copied)
        self._taken = idx
        return slot.views

    def release(self) -> None:
        """mark the taken set as consumed at this point of the current stream"""
        if self._taken is None:
            raise RuntimeError("nothing taken")
        slot = self._slots[self._taken]
        slot.consumed.record(torch.cuda.current_stream(self.device))
        slot.used = True

    def drain(self) -> None:
        """make the current stream wait for every copy issued so far"""
        torch.cuda.current_stream(self.device).wait_stream(self.copy_stream)
