"""One staging buffer: a drain blocks until the write finished (reference: ``io/single_io_buffer.py``)."""
from __future__ import annotations

from .base_io_buffer import NUM_BYTES_ALIGNED, Base_IO_Buffer


class Single_IO_Buffer(Base_IO_Buffer):

    def __init__(self, pinned_tensor, dnvme_handle):
        super().__init__(pinned_tensor, dnvme_handle)
        self._pinned_offset = 0

    def fill(self, src_tensor, src_offset):
        n = Base_IO_Buffer.fill_buffer(src_tensor, src_offset, self._pinned_tensor, self._pinned_offset)
        self._pinned_offset += n
        return n

    def drain(self, num_bytes, fd, file_offset):
        self._drain(num_bytes, fd, file_offset, blocking=True)
        tail = self._pinned_offset - num_bytes
        if tail:        # carry the unaligned remainder to the front
            self._pinned_tensor[:tail].copy_(self._pinned_tensor[num_bytes:self._pinned_offset].clone())
        self._pinned_offset = tail

    def get_buffer(self): return self._pinned_tensor
    def get_offset(self): return self._pinned_offset
    def get_aligned_num_bytes(self): return self._pinned_offset // NUM_BYTES_ALIGNED * NUM_BYTES_ALIGNED
    def get_unaligned_num_bytes(self): return self._pinned_offset % NUM_BYTES_ALIGNED
    def is_full(self): return self._pinned_offset == self._pinned_tensor.numel()
    def is_empty(self): return self._pinned_offset == 0
    def reset(self): self._pinned_offset = 0
