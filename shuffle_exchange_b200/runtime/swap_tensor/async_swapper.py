"""Write-behind staging for a stream of tensors headed to swap files.

A small set of pinned *slots* is cycled through three states::

    free  --claim-->  filling  --submit-->  in flight  --drain-->  free

Tensors are copied back to back (aligned for O_DIRECT) into the slot that is *filling*.  When the next tensor does not
fit, the slot's writes are submitted to the IO engine (``csrc/host/aio.cpp`` queue through the ``aio_handle``) and the
next free slot is claimed; only when no slot is free does the caller block, on the engine's ``wait()``, which drains
everything in flight at once.  ``flush()`` submits the partial slot and drains.  Used for gradient swap-out by the
NVMe optimizer swappers (the reference's equivalent: ``runtime/swap_tensor/async_swapper.py``, same role and the same
``add_buffers / swap_out_tensors / release_buffers`` surface)."""
from __future__ import annotations

from collections import deque
from typing import Deque, List, Optional, Tuple

import torch

ASYNC_SWAPPER_WAIT_TIMER = "async_swap_gradient_wait"
INVALID_BUFFER_INDEX = -1


class _Slot:
    """One pinned staging buffer with a bump pointer and the writes it will issue."""

    def __init__(self, buffer: torch.Tensor):
        self.buffer = buffer
        self.used = 0
        self.writes: List[Tuple[torch.Tensor, str]] = []       # (aligned view, file)

    def room_for(self, numel: int) -> bool:
        return self.used + numel <= self.buffer.numel()

    def stage(self, tensor: torch.Tensor, path: str, aligned: int) -> None:
        view = self.buffer.narrow(0, self.used, aligned)
        view.narrow(0, 0, tensor.numel()).copy_(tensor.reshape(-1))
        self.writes.append((view, path))
        self.used += aligned

    def clear(self) -> None:
        self.used = 0
        self.writes = []


class AsyncTensorSwapper:
    def __init__(self, aio_handle, numel_alignment, timers=None):
        self.aio_handle = aio_handle
        self.numel_alignment = int(numel_alignment)
        self.timers = timers
        self.timer_names = set()
        self._slots: List[_Slot] = []
        self._free: Deque[int] = deque()
        self._filling: Optional[int] = None
        self._in_flight: List[int] = []
        self._pending_writes = 0
        self.num_elements_swapped = 0
        self.max_numel = 0
        self.dtype = None

    # ------------------------------------------------------------------------------- slots
    def has_buffers(self) -> bool:
        return bool(self._slots)

    def add_buffers(self, buffer_list) -> None:
        if self._slots:
            raise RuntimeError("staging buffers are already attached; release_buffers() first")
        kinds = {(b.device.type, b.dtype) for b in buffer_list}
        if len(kinds) != 1 or next(iter(kinds))[0] != "cpu":
            raise ValueError("staging buffers must be host tensors of one dtype")
        self.dtype = buffer_list[0].dtype
        self._slots = [_Slot(b) for b in buffer_list]
        self._free = deque(range(len(self._slots)))
        self.max_numel = max(b.numel() for b in buffer_list)

    def release_buffers(self):
        """Drain everything and hand the pinned buffers back to the caller."""
        self.flush()
        buffers = [s.buffer for s in self._slots]
        self._slots, self._free, self._filling = [], deque(), None
        self.num_elements_swapped = 0
        self.dtype = None
        return buffers

    def get_timer_names(self):
        return sorted(self.timer_names)

    # ------------------------------------------------------------------------------ writes
    def _aligned(self, numel: int) -> int:
        a = self.numel_alignment
        return (numel + a - 1) // a * a

    def swap_out_tensors(self, tensor_list, path_list) -> None:
        if not self._slots:
            raise RuntimeError("add_buffers() must be called before swap_out_tensors()")
        for tensor, path in zip(tensor_list, path_list):
            need = self._aligned(tensor.numel())
            if need > self.max_numel:
                raise ValueError(f"tensor of {tensor.numel()} elements does not fit a staging buffer of {self.max_numel}")
            if self._filling is not None and not self._slots[self._filling].room_for(need):
                self._submit_filling()
            if self._filling is None:
                if not self._free:
                    self._drain()
                self._filling = self._free.popleft()
            self._slots[self._filling].stage(tensor, path, need)

    def _submit_filling(self) -> None:
        idx, self._filling = self._filling, None
        if idx is None:
            return
        slot = self._slots[idx]
        for view, path in slot.writes:
            if self.aio_handle.async_pwrite(view, path) != 0:
                raise IOError(f"async write to {path} could not be queued")
        self._pending_writes += len(slot.writes)
        self._in_flight.append(idx)

    def _drain(self) -> None:
        if self._pending_writes:
            timer = self.timers(ASYNC_SWAPPER_WAIT_TIMER) if self.timers else None
            if timer:
                self.timer_names.add(ASYNC_SWAPPER_WAIT_TIMER)
                timer.start()
            completed = self.aio_handle.wait()
            if timer:
                timer.stop()
            if completed != self._pending_writes:
                raise IOError(f"IO engine completed {completed} of {self._pending_writes} queued writes")
            self._pending_writes = 0
        for idx in self._in_flight:
            self.num_elements_swapped += self._slots[idx].used
            self._slots[idx].clear()
            self._free.append(idx)
        self._in_flight = []

    def flush(self) -> None:
        """Submit the partially filled slot and wait until every queued write has completed."""
        self._submit_filling()
        self._drain()

    # name the optimizer swappers were written against
    _flush_buffers_until_complete = flush
