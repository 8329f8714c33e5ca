"""Device-tensor file IO handle (the ``GDSBuilder().load().gds_handle`` of the reference, ``ops/gds`` + the absent
``csrc/gds`` cuFile binding).

``gds_handle`` has the ``aio_handle`` method surface but accepts CUDA tensors.  cuFile is not part of this image, so
transfers take the path cuFile itself uses in compatibility mode: the tensor is streamed through two pinned bounce
buffers on a copy stream (DMA engine, not SMs) while the native async-IO engine (``csrc/host/aio.cpp``) reads or writes
the other buffer — device<->host and host<->NVMe overlap chunk by chunk.  ``pin_device_tensor`` keeps the
registration bookkeeping the reference API exposes."""
from __future__ import annotations

import sys
from typing import Dict

import torch

from ..aio import aio_handle


class gds_handle(aio_handle):
    BOUNCE_BYTES = 32 << 20

    def __init__(self, block_size: int = 1 << 20, queue_depth: int = 32, single_submit: bool = False,
                 overlap_events: bool = True, intra_op_parallelism: int = 4):
        super().__init__(block_size, queue_depth, single_submit, overlap_events, intra_op_parallelism)
        self._registered: Dict[int, torch.Tensor] = {}
        self._bounce = None
        self._stream = None

    def _buffers(self):
        if self._bounce is None:
            pin = torch.cuda.is_available()
            self._bounce = [torch.empty(self.BOUNCE_BYTES, dtype=torch.uint8, pin_memory=pin) for _ in range(2)]
            self._stream = torch.cuda.Stream() if pin else None
        return self._bounce

    # ------------------------------------------------------------------------------------------ registration
    def pin_device_tensor(self, t: torch.Tensor):
        self._registered[t.data_ptr()] = t
        return t

    def unpin_device_tensor(self, t: torch.Tensor):
        self._registered.pop(t.data_ptr(), None)

    def new_pinned_device_tensor(self, num_elem: int, like: torch.Tensor) -> torch.Tensor:
        return self.pin_device_tensor(torch.empty(num_elem, dtype=like.dtype, device=like.device))

    def free_pinned_device_tensor(self, t: torch.Tensor) -> None:
        self.unpin_device_tensor(t)

    # ------------------------------------------------------------------------------------------ IO
    def _io(self, write: bool, buffer: torch.Tensor, filename: str, is_async: bool, file_offset: int = 0) -> int:
        if buffer.device.type == "cpu":
            return super()._io(write, buffer, filename, is_async, file_offset)
        assert buffer.is_contiguous(), "gds works on contiguous device tensors"
        flat = buffer.view(-1).view(torch.uint8)
        total = flat.numel()
        bufs = self._buffers()
        cur = torch.cuda.current_stream()
        self._stream.wait_stream(cur)
        step = self.BOUNCE_BYTES
        pending = [None, None]          # per bounce buffer: (offset, nbytes) of an IO in flight / data staged
        done = 0
        i = 0
        with torch.cuda.stream(self._stream):
            while done < total:
                n = min(step, total - done)
                b = bufs[i & 1][:n]
                if write:
                    b.copy_(flat[done:done + n], non_blocking=True)
                    self._stream.synchronize()                       # chunk is in host memory
                    super().wait()                                   # previous chunk's write finished: overlap of 1
                    super()._io(True, b, filename, True, file_offset + done)
                else:
                    super()._io(False, b, filename, False, file_offset + done)
                    flat[done:done + n].copy_(b, non_blocking=True)
                    self._stream.synchronize()
                done += n
                i += 1
            if write:
                super().wait()
        cur.wait_stream(self._stream)
        return 0

    def wait(self) -> int:
        return super().wait()


class GDSBuilder:
    """Builder-shaped accessor (``GDSBuilder().load().gds_handle(...)``)."""
    NAME = "gds"

    def is_compatible(self, verbose=False):
        return True

    def load(self):
        return sys.modules[__name__]
