"""Asynchronous file IO handle (the ``AsyncIOBuilder().load().aio_handle`` of the reference).

Native engine: ``csrc/host/aio.cpp`` (thread pool issuing positional reads/writes, ``O_DIRECT`` when aligned).  The
method surface matches what the swap-tensor and checkpoint code of the reference call
(``runtime/swap_tensor/utils.py:22-27``, ``partitioned_param_swapper.py:111-122``, ``io/fast_file_writer.py``):
``sync_pread / sync_pwrite / async_pread / async_pwrite / pread / pwrite / wait`` and the pinned-buffer helpers.
"""
from __future__ import annotations

import ctypes
import os

import torch

from ...native import host as _host

c_p, c_i, c_l = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong
_host.register({
    "sxbh_aio_create": [c_l, c_i, c_i, c_i, c_i],
    "sxbh_aio_destroy": [c_i],
    "sxbh_aio_pread": [c_i, c_p, c_l, ctypes.c_char_p, c_l, c_i],
    "sxbh_aio_pwrite": [c_i, c_p, c_l, ctypes.c_char_p, c_l, c_i],
    "sxbh_aio_wait": [c_i],
    "sxbh_aio_info": [c_i, c_i],
    "sxbh_file_size": [ctypes.c_char_p],
})


class aio_handle:
    def __init__(self, block_size: int = 1 << 20, queue_depth: int = 32, single_submit: bool = False,
                 overlap_events: bool = True, intra_op_parallelism: int = 4):
        self._lib = _host.load()
        self._lib.sxbh_aio_info.restype = c_l
        self._lib.sxbh_file_size.restype = c_l
        self._id = self._lib.sxbh_aio_create(int(block_size), int(queue_depth), int(single_submit), int(overlap_events),
                                             int(intra_op_parallelism))
        self._keepalive = []          # tensors of in-flight async requests
        self._pinned = {}

    def __del__(self):
        try:
            self._lib.sxbh_aio_destroy(self._id)
        except Exception:
            pass

    # configuration getters (reference names)
    def get_block_size(self): return int(self._lib.sxbh_aio_info(self._id, 0))
    def get_queue_depth(self): return int(self._lib.sxbh_aio_info(self._id, 1))
    def get_single_submit(self): return bool(self._lib.sxbh_aio_info(self._id, 2))
    def get_overlap_events(self): return bool(self._lib.sxbh_aio_info(self._id, 3))
    def get_intra_op_parallelism(self): return int(self._lib.sxbh_aio_info(self._id, 4))
    def get_thread_count(self): return self.get_intra_op_parallelism()

    def _io(self, write: bool, buffer: torch.Tensor, filename: str, is_async: bool, file_offset: int = 0) -> int:
        assert buffer.device.type == "cpu" and buffer.is_contiguous(), "aio works on contiguous host tensors"
        nbytes = buffer.numel() * buffer.element_size()
        fn = self._lib.sxbh_aio_pwrite if write else self._lib.sxbh_aio_pread
        rc = fn(self._id, buffer.data_ptr(), nbytes, os.fsencode(filename), int(file_offset), int(is_async))
        if rc < 0:
            raise OSError(-rc, f"aio {'write' if write else 'read'} of {filename} failed: {os.strerror(-rc)}")
        if is_async:
            self._keepalive.append(buffer)
        return 0

    def sync_pread(self, buffer, filename, file_offset=0): return self._io(False, buffer, filename, False, file_offset)
    def sync_pwrite(self, buffer, filename, file_offset=0): return self._io(True, buffer, filename, False, file_offset)
    def async_pread(self, buffer, filename, file_offset=0): return self._io(False, buffer, filename, True, file_offset)
    def async_pwrite(self, buffer, filename, file_offset=0): return self._io(True, buffer, filename, True, file_offset)

    def pread(self, buffer, filename, validate=False, async_op=False, file_offset=0):
        return self._io(False, buffer, filename, async_op, file_offset)

    def pwrite(self, buffer, filename, validate=False, async_op=False, file_offset=0):
        return self._io(True, buffer, filename, async_op, file_offset)

    def read(self, buffer, filename, validate=False): return self.sync_pread(buffer, filename)
    def write(self, buffer, filename, validate=False): return self.sync_pwrite(buffer, filename)

    def wait(self) -> int:
        rc = self._lib.sxbh_aio_wait(self._id)
        self._keepalive.clear()
        if rc < 0:
            raise OSError(-rc, f"aio wait failed: {os.strerror(-rc)}")
        return rc

    # pinned staging buffers ------------------------------------------------------------------------
    def new_cpu_locked_tensor(self, num_elem: int, like: torch.Tensor) -> torch.Tensor:
        """Page-locked host tensor, 4 KiB aligned (cudaHostAlloc is page aligned) so IO goes O_DIRECT."""
        t = torch.empty(num_elem, dtype=like.dtype, device="cpu", pin_memory=torch.cuda.is_available())
        self._pinned[t.data_ptr()] = t
        return t

    def free_cpu_locked_tensor(self, t: torch.Tensor) -> None:
        self._pinned.pop(t.data_ptr(), None)

    def pin_device_tensor(self, t): return t
    def unpin_device_tensor(self, t): return None


def file_size(path: str) -> int:
    lib = _host.load()
    lib.sxbh_file_size.restype = c_l
    return int(lib.sxbh_file_size(os.fsencode(path)))


class AsyncIOBuilder:
    """Builder-shaped accessor kept for call-site compatibility (``AsyncIOBuilder().load().aio_handle(...)``)."""
    NAME = "async_io"

    def is_compatible(self, verbose=False):
        return True

    def load(self):
        import sys
        return sys.modules[__name__]
