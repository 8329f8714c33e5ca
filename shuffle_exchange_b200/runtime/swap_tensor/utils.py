"""Building blocks of tensor swapping: batched async reads/writes and pinned staging buffers carved into aligned
slots (reference: ``runtime/swap_tensor/utils.py:18-240``)."""
from __future__ import annotations

from typing import List

import torch

from ...utils.logging import logger

MIN_AIO_BYTES = 1024 ** 2
AIO_ALIGNED_BYTES = 1024


def swap_in_tensors(swap_handle, tensor_buffers: List[torch.Tensor], swap_paths: List[str]) -> None:
    for buf, path in zip(tensor_buffers, swap_paths):
        assert swap_handle.async_pread(buf, path) == 0


def swap_out_tensors(swap_handle, tensor_buffers: List[torch.Tensor], swap_paths: List[str]) -> None:
    for buf, path in zip(tensor_buffers, swap_paths):
        assert swap_handle.async_pwrite(buf, path) == 0


def print_object(obj, name, exclude_list=[]):
    logger.info("{}:".format(name))
    for arg in sorted(vars(obj)):
        if arg not in exclude_list:
            logger.info("  {} {} {}".format(arg, "." * (29 - len(arg)), getattr(obj, arg)))


def get_sized_buffer(buffer, num_elems):
    assert num_elems <= buffer.numel(), f"num_elems {num_elems} > buffer {buffer.numel()}"
    return buffer.narrow(0, 0, num_elems) if num_elems < buffer.numel() else buffer


def get_sized_buffers(buffer_list, num_elems_list):
    return [get_sized_buffer(b, n) for b, n in zip(buffer_list, num_elems_list)]


class SwapBuffer:
    """One pinned buffer filled front to back with tensors destined for (or read from) swap files; every tensor starts
    at an aligned offset and occupies an aligned length so that IO can go ``O_DIRECT``."""

    def __init__(self, buffer: torch.Tensor):
        self.buffer = buffer
        self.reset()

    def reset(self):
        self.offset = 0
        self.swap_tensors, self.compute_tensors, self.swap_paths = {}, {}, {}
        self.num_elem = 0

    def insert_tensor(self, tensor, swap_path, aligned_numel):
        swap_t, compute_t = self.allocate_tensor(swap_path, tensor.numel(), aligned_numel)
        compute_t.data.copy_(tensor.data.reshape(-1))
        return swap_t, compute_t

    def allocate_tensor(self, swap_path, numel, aligned_numel):
        assert self.has_space(aligned_numel)
        assert self.offset not in self.swap_tensors
        swap_t = self.buffer.narrow(0, self.offset, aligned_numel)
        compute_t = swap_t.narrow(0, 0, numel)
        self.swap_tensors[self.offset], self.compute_tensors[self.offset] = swap_t, compute_t
        self.swap_paths[self.offset] = swap_path
        self.offset += aligned_numel
        self.num_elem += numel
        return swap_t, compute_t

    def has_space(self, numel):
        return self.offset + numel <= self.buffer.numel()

    def get_swap_tensors(self):
        return list(self.swap_tensors.values())

    def get_swap_paths(self):
        return list(self.swap_paths.values())

    def get_compute_tensors(self):
        return list(self.compute_tensors.values())

    def get_num_elem(self):
        return self.num_elem

    def get_swap_tensor(self, offset):
        return self.swap_tensors.get(offset)

    def get_compute_tensor(self, offset):
        return self.compute_tensors.get(offset)

    def get_swap_path(self, offset):
        return self.swap_paths.get(offset)


class SwapBufferPool:
    def __init__(self, buffers: List[torch.Tensor]):
        assert all(b.device.type == "cpu" for b in buffers)
        self.buffers = [SwapBuffer(b) for b in buffers]
        self.current_index = 0

    def reset(self):
        self.current_index = 0
        for b in self.buffers:
            b.reset()

    def _current(self):
        return self.buffers[self.current_index]

    def allocate_tensor(self, numel, swap_path, aligned_numel):
        if self.has_space(aligned_numel):
            return self._current().allocate_tensor(swap_path, numel, aligned_numel)
        return None, None

    def insert_tensor(self, tensor, swap_path, aligned_numel):
        if self.has_space(aligned_numel):
            return self._current().insert_tensor(tensor, swap_path, aligned_numel)
        return None, None

    def get_swap_tensors(self):
        return [t for b in self._used() for t in b.get_swap_tensors()]

    def get_swap_paths(self):
        return [p for b in self._used() for p in b.get_swap_paths()]

    def get_compute_tensors(self):
        return [t for b in self._used() for t in b.get_compute_tensors()]

    def has_space(self, numel):
        if self._current().has_space(numel):
            return True
        if self.current_index == len(self.buffers) - 1:
            return False
        self.current_index += 1
        return self._current().has_space(numel)

    def swap_out(self, aio_handle, async_op=False):
        tensors, paths = self.get_swap_tensors(), self.get_swap_paths()
        assert all(p is not None for p in paths)
        swap_out_tensors(aio_handle, tensors, paths)
        if not async_op:
            assert len(tensors) == aio_handle.wait()

    def swap_in(self, aio_handle, async_op=False):
        tensors, paths = self.get_swap_tensors(), self.get_swap_paths()
        assert all(p is not None for p in paths)
        swap_in_tensors(aio_handle, tensors, paths)
        if not async_op:
            assert len(tensors) == aio_handle.wait()

    def _used(self):
        return self.buffers[: self.current_index + 1]


class SwapBufferManager:
    """``count`` pinned buffers of ``num_elems`` elements handed out and returned by id."""

    def __init__(self, num_elems, count, dtype):
        self.num_elems, self.count, self.dtype = num_elems, count, dtype
        pin = torch.cuda.is_available()
        self.all_buffers = [torch.zeros(num_elems, device="cpu", dtype=dtype, pin_memory=pin) for _ in range(count)]
        self.free_buffer_index = list(range(count))
        self.used_buffer_index = {}
        self.gigabytes = (self.all_buffers[0].element_size() * num_elems * count) / (1024 ** 3)

    def allocate(self, num_elems, count, dtype):
        assert dtype == self.dtype and num_elems <= self.num_elems
        if count > len(self.free_buffer_index):
            return None
        picked = self.free_buffer_index[-count:]
        self.free_buffer_index = self.free_buffer_index[:-count]
        out = []
        for i in picked:
            t = self.all_buffers[i].narrow(0, 0, num_elems)
            out.append(t)
            self.used_buffer_index[id(t)] = i
        return out

    def allocate_all(self, num_elems, dtype):
        return self.allocate(num_elems=num_elems, count=len(self.free_buffer_index), dtype=dtype)

    def free(self, buffers):
        ids = [id(t) for t in buffers]
        assert all(i in self.used_buffer_index for i in ids)
        for i in ids:
            self.free_buffer_index.append(self.used_buffer_index.pop(i))
