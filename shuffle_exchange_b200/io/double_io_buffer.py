"""Two half-buffers: one fills while the other drains asynchronously (reference: ``io/double_io_buffer.py``)."""
from __future__ import annotations

from .base_io_buffer import NUM_BYTES_ALIGNED, Base_IO_Buffer

INVALID_BUFFER_INDEX = -1


class Double_IO_Buffer(Base_IO_Buffer):

    def __init__(self, pinned_tensor, dnvme_handle):
        super().__init__(pinned_tensor, dnvme_handle)
        assert pinned_tensor.numel() % (2 * NUM_BYTES_ALIGNED) == 0
        half = pinned_tensor.numel() // 2
        self._buffers = [pinned_tensor[:half], pinned_tensor[half:]]
        self._buffer_index = 0                      # the half being filled
        self._drain_index = INVALID_BUFFER_INDEX    # the half with a write in flight
        self._buffer_offset = 0

    def fill(self, src_tensor, src_offset):
        n = Base_IO_Buffer.fill_buffer(src_tensor, src_offset, self._buffers[self._buffer_index], self._buffer_offset)
        self._buffer_offset += n
        return n

    def drain(self, num_bytes, fd, file_offset):
        self._validate_buffer_index(self._buffer_index)
        self.complete_ongoing_drain()               # at most one write in flight
        self._drain(num_bytes, fd, file_offset, blocking=False)
        self._drain_index = self._buffer_index
        nxt = self._buffer_index ^ 1
        tail = self._buffer_offset - num_bytes
        if tail:                                    # unaligned remainder moves to the front of the other half
            self._buffers[nxt][:tail].copy_(self._buffers[self._buffer_index][num_bytes:self._buffer_offset])
        self._buffer_index, self._buffer_offset = nxt, tail

    def complete_ongoing_drain(self):
        if self._is_ongoing_drain():
            self._wait_for_drain()

    def get_buffer(self):
        self._validate_buffer_index(self._buffer_index)
        return self._buffers[self._buffer_index]

    def get_offset(self): return self._buffer_offset
    def get_aligned_num_bytes(self): return self._buffer_offset // NUM_BYTES_ALIGNED * NUM_BYTES_ALIGNED
    def get_unaligned_num_bytes(self): return self._buffer_offset % NUM_BYTES_ALIGNED
    def is_full(self): return self._buffer_offset == self._buffers[self._buffer_index].numel()
    def is_empty(self): return self._buffer_offset == 0 and not self._is_ongoing_drain()

    def reset(self):
        self.complete_ongoing_drain()
        self._buffer_index, self._buffer_offset = 0, 0

    def _validate_buffer_index(self, index):
        assert index in (0, 1), f"invalid buffer index {index}"

    def _wait_for_drain(self):
        self._validate_buffer_index(self._drain_index)
        self._dnvme_handle.wait()
        self._drain_index = INVALID_BUFFER_INDEX

    def _is_ongoing_drain(self):
        return self._drain_index != INVALID_BUFFER_INDEX
