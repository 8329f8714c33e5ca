"""Pinned staging buffers between tensors and the asynchronous file engine (reference: ``io/base_io_buffer.py``).

A buffer is a pinned byte tensor that ``fill`` copies tensor bytes into and ``drain`` writes to a file through an aio
handle.  Direct IO wants sector-aligned lengths, so a drain writes the aligned prefix and carries the unaligned tail to
the front of the (next) buffer; ``get_unaligned_num_bytes`` is what a final buffered write must flush."""
from __future__ import annotations

import torch

NUM_BYTES_ALIGNED = 1024


class Base_IO_Buffer:

    def __init__(self, pinned_tensor: torch.Tensor, dnvme_handle):
        assert pinned_tensor.dtype == torch.uint8 and pinned_tensor.dim() == 1, "staging buffer must be 1-D uint8"
        assert pinned_tensor.numel() % NUM_BYTES_ALIGNED == 0, f"buffer size must be a multiple of {NUM_BYTES_ALIGNED}"
        self._dnvme_handle = dnvme_handle
        self._pinned_tensor = pinned_tensor

    # ---- interface
    def fill(self, src_tensor, src_offset): raise NotImplementedError
    def drain(self, num_bytes, fd, file_offset): raise NotImplementedError
    def is_empty(self): raise NotImplementedError
    def is_full(self): raise NotImplementedError
    def get_buffer(self): raise NotImplementedError
    def get_offset(self): raise NotImplementedError
    def get_aligned_num_bytes(self): raise NotImplementedError
    def get_unaligned_num_bytes(self): raise NotImplementedError
    def reset(self): raise NotImplementedError

    def complete_ongoing_drain(self):
        pass

    # ---- shared mechanics
    def _drain(self, num_bytes: int, fd, file_offset: int, blocking: bool = False):
        """Write the first ``num_bytes`` (aligned) of the active buffer at ``file_offset`` of ``fd`` (a path)."""
        assert num_bytes <= self.get_offset() and num_bytes % NUM_BYTES_ALIGNED == 0
        buf = self.get_buffer()[:num_bytes]
        if blocking:
            self._dnvme_handle.sync_pwrite(buf, fd, file_offset)
        else:
            self._dnvme_handle.async_pwrite(buf, fd, file_offset)

    @staticmethod
    def fill_buffer(src_tensor: torch.Tensor, src_offset: int, buffer_tensor: torch.Tensor, buffer_offset: int) -> int:
        """Copy as many bytes of ``src_tensor`` (flat uint8 view), starting at ``src_offset``, as fit behind
        ``buffer_offset``; returns the count."""
        src = src_tensor.reshape(-1).view(torch.uint8) if src_tensor.dtype != torch.uint8 else src_tensor.reshape(-1)
        n = min(src.numel() - src_offset, buffer_tensor.numel() - buffer_offset)
        if n > 0:
            buffer_tensor.narrow(0, buffer_offset, n).copy_(src.narrow(0, src_offset, n), non_blocking=False)
        return max(n, 0)
