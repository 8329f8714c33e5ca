"""Fire-and-forget swap-out of a stream of tensors through a small ring of pinned buffers (reference:
``runtime/swap_tensor/async_swapper.py:19-175``): tensors are packed into the current buffer; when it is full it is
written asynchronously while packing continues in the next one."""
from __future__ import annotations

import torch

from ...utils.logging import logger
from .utils import SwapBufferPool, get_sized_buffer, swap_out_tensors

INVALID_BUFFER_INDEX = -1
ASYNC_SWAPPER_WAIT_TIMER = "async_swap_gradient_wait"


class AsyncTensorSwapper:
    def __init__(self, aio_handle, numel_alignment, timers=None):
        self.free_buffer_index, self.swapping_buffer_index, self.ready_buffer_index = [], [], []
        self.current_buffer_index = INVALID_BUFFER_INDEX
        self.all_buffers = []
        self.aio_handle = aio_handle
        self.numel_alignment = numel_alignment
        self.max_numel = 0
        self.num_pending_swaps = 0
        self.timers = timers
        self.timer_names = set()
        self.num_elements_swapped = 0
        self.dtype = None

    def has_buffers(self):
        return len(self.all_buffers) > 0

    def add_buffers(self, buffer_list):
        assert len(self.all_buffers) == 0 and all(b.device.type == "cpu" for b in buffer_list)
        dtype = buffer_list[0].dtype
        assert all(b.dtype == dtype for b in buffer_list)
        self.dtype = dtype
        self.all_buffers = [SwapBufferPool([b]).buffers[0] for b in buffer_list]
        self.free_buffer_index += list(range(len(self.all_buffers)))
        self.max_numel = max(b.numel() for b in buffer_list)

    def get_timer_names(self):
        return list(self.timer_names)

    def release_buffers(self):
        self._flush_buffers_until_complete()
        pinned = [b.buffer for b in self.all_buffers]
        self.all_buffers, self.free_buffer_index = [], []
        self.current_buffer_index = INVALID_BUFFER_INDEX
        self.num_elements_swapped = 0
        self.dtype = None
        return pinned

    def swap_out_tensors(self, tensor_list, path_list):
        for t, p in zip(tensor_list, path_list):
            self._swap_out_tensor(t, p)

    def _aligned(self, numel):
        r = numel % self.numel_alignment
        return numel if r == 0 else numel + self.numel_alignment - r

    def _swap_out_tensor(self, tensor, swap_path):
        assert len(self.all_buffers) > 0
        aligned = self._aligned(tensor.numel())
        assert aligned <= self.max_numel
        self._make_swap_space(aligned)
        assert self.current_buffer_index != INVALID_BUFFER_INDEX
        self.all_buffers[self.current_buffer_index].insert_tensor(tensor, swap_path, aligned)

    def _make_swap_space(self, numel):
        if self.current_buffer_index == INVALID_BUFFER_INDEX:
            self._allocate_buffer()
            return
        if not self.all_buffers[self.current_buffer_index].has_space(numel):
            if len(self.free_buffer_index) > 0:
                self._flush_ready_buffers()
            else:
                self._flush_buffers_until_complete()
            self._allocate_buffer()

    def _allocate_buffer(self):
        assert self.current_buffer_index == INVALID_BUFFER_INDEX and len(self.free_buffer_index) > 0
        self.current_buffer_index = self.free_buffer_index.pop()
        self.all_buffers[self.current_buffer_index].reset()

    def _flush_ready_buffers(self):
        if self.current_buffer_index != INVALID_BUFFER_INDEX:
            self.ready_buffer_index.append(self.current_buffer_index)
            self.current_buffer_index = INVALID_BUFFER_INDEX
        self._swap_out_ready_buffers()

    def _flush_buffers_until_complete(self):
        self._flush_ready_buffers()
        assert len(self.ready_buffer_index) == 0
        self._wait_for_swap_complete()
        assert len(self.swapping_buffer_index) == 0 and len(self.free_buffer_index) == len(self.all_buffers)

    def _swap_out_ready_buffers(self):
        for idx in self.ready_buffer_index:
            buf = self.all_buffers[idx]
            tensors, paths = buf.get_swap_tensors(), buf.get_swap_paths()
            self.num_pending_swaps += len(tensors)
            swap_out_tensors(self.aio_handle, tensors, paths)
        self.swapping_buffer_index += self.ready_buffer_index
        self.ready_buffer_index = []

    def _wait_for_swap_complete(self):
        assert len(self.swapping_buffer_index) > 0 or self.num_pending_swaps == 0
        if self.num_pending_swaps:
            assert self.aio_handle.wait() == self.num_pending_swaps
        self.num_pending_swaps = 0
        for idx in self.swapping_buffer_index:
            self.num_elements_swapped += self.all_buffers[idx].get_num_elem()
            self.all_buffers[idx].reset()
        self.free_buffer_index += self.swapping_buffer_index
        self.swapping_buffer_index = []
