"""``FastFileWriter``: file-like sink that packs everything ``torch.save`` emits into large pinned buffers and writes
them with the asynchronous IO engine, double-buffered so that serialisation of chunk *k+1* overlaps the NVMe write of
chunk *k*.

Reference: ``deepspeed/io/fast_file_writer.py:35-262`` (+ ``{single,double}_io_buffer.py``).  Differences: the two
buffers and the in-flight bookkeeping live in this one class; the last, unaligned tail of the file is written through a
plain positional write instead of a second aligned pass + truncate.
"""
from __future__ import annotations

import os
import time
import warnings
from dataclasses import dataclass
from typing import Optional

import torch

from ..ops.aio import aio_handle
from .base_file_writer import BaseFileWriter
from .constants import *  # noqa: F401,F403

warnings.filterwarnings("ignore", message="The given buffer is not writable")   # read-only views of torch.save's bytes

FASTIO_STAT_KEYS = [AIO_WRITE_SEC_KEY, AIO_WRITE_BYTES_KEY, AIO_SPEED_KEY, SLOW_WRITE_BYTES_KEY, SLOW_WRITE_SEC_KEY,
                    AIO_FILL_BUFFER_COUNT_KEY, AIO_FILL_BUFFER_SEC_KEY, AIO_FILL_BUFFER_SPEED_KEY]


@dataclass
class FastFileWriterConfig:
    dnvme_handle: object = None
    pinned_tensor: Optional[torch.Tensor] = None        # uint8 staging memory (split in two halves when double_buffer)
    double_buffer: bool = True
    num_parallel_writers: int = 1
    writer_rank: int = 0
    global_rank: int = 0


class FastFileWriter(BaseFileWriter):
    ALIGN = 4096

    def __init__(self, file_path, config: FastFileWriterConfig = None, buffer_bytes: int = 64 << 20):
        super().__init__(file_path)
        config = config or FastFileWriterConfig()
        self._handle = config.dnvme_handle or aio_handle(1 << 20, 32, False, True, 4)
        pinned = config.pinned_tensor
        if pinned is None:
            pinned = torch.empty(buffer_bytes, dtype=torch.uint8, pin_memory=torch.cuda.is_available())
        assert pinned.dtype == torch.uint8 and pinned.device.type == "cpu"
        n = pinned.numel()
        if config.double_buffer:
            half = (n // 2) // self.ALIGN * self.ALIGN
            self._bufs = [pinned[:half], pinned[half: 2 * half]]
        else:
            self._bufs = [pinned[: n // self.ALIGN * self.ALIGN]]
        self._cur, self._fill = 0, 0
        self._file_offset = 0
        self._inflight = [False] * len(self._bufs)
        for k in FASTIO_STAT_KEYS:
            self._stats[k] = 0
        open(file_path, "wb").close()                      # create / truncate

    # ------------------------------------------------------------------------------ file-like API
    def write(self, buffer) -> int:
        t0 = time.time()
        mv = memoryview(buffer).cast("B")
        n = len(mv)
        src = torch.frombuffer(mv, dtype=torch.uint8) if n else None
        done = 0
        while done < n:
            buf = self._bufs[self._cur]
            room = buf.numel() - self._fill
            take = min(room, n - done)
            t1 = time.time()
            buf[self._fill: self._fill + take].copy_(src[done: done + take])
            self._incr_stats(AIO_FILL_BUFFER_SEC_KEY, time.time() - t1)
            self._incr_stats(AIO_FILL_BUFFER_COUNT_KEY)
            self._fill += take
            done += take
            if self._fill == buf.numel():
                self._submit_current(full=True)
        self._incr_stats(WRITE_COUNT_KEY)
        self._incr_stats(WRITE_BYTES_KEY, n)
        self._incr_stats(WRITE_SEC_KEY, time.time() - t0)
        return n

    def save_torch_storage_object_list(self, storage_obj_list, save_size):
        """Bulk path for (storage, dtype) pairs: bytes go straight into the staging buffers."""
        t0 = time.time()
        total = 0
        for storage in storage_obj_list:
            t = torch.frombuffer(storage.untyped(), dtype=torch.uint8) if not torch.is_tensor(storage) else storage.view(torch.uint8)
            total += self.write(t.numpy().data)
        self._incr_stats(SAVE_STORAGE_KEY, len(storage_obj_list))
        self._incr_stats(SAVE_STORAGE_BYTES_KEY, total)
        self._incr_stats(SAVE_STORAGE_SEC_KEY, time.time() - t0)
        return total

    def flush(self):
        self._incr_stats(FLUSH_COUNT_KEY)

    def fileno(self):
        return INVALID_FD

    def close(self):
        self._finish()
        self._incr_stats(CLOSE_COUNT_KEY)

    def __del__(self):
        try:
            self._finish()
        except Exception:
            pass

    # ---------------------------------------------------------------------------------- internals
    def _wait_all(self):
        if any(self._inflight):
            t0 = time.time()
            self._handle.wait()
            self._incr_stats(AIO_WRITE_SEC_KEY, time.time() - t0)
            self._inflight = [False] * len(self._bufs)

    def _submit_current(self, full: bool):
        buf = self._bufs[self._cur]
        nbytes = self._fill if not full else buf.numel()
        aligned = nbytes // self.ALIGN * self.ALIGN
        if len(self._bufs) == 1:
            self._wait_all()
        if aligned:
            self._handle.async_pwrite(buf[:aligned], self._file_path, file_offset=self._file_offset)
            self._inflight[self._cur] = True
            self._incr_stats(AIO_WRITE_BYTES_KEY, aligned)
        tail = nbytes - aligned
        if tail:                                  # only at close: unaligned remainder through a positional write
            self._wait_all()
            t0 = time.time()
            with open(self._file_path, "r+b") as f:
                f.seek(self._file_offset + aligned)
                f.write(buf[aligned:nbytes].numpy().tobytes())
            self._incr_stats(SLOW_WRITE_BYTES_KEY, tail)
            self._incr_stats(SLOW_WRITE_SEC_KEY, time.time() - t0)
        self._file_offset += nbytes
        self._fill = 0
        if len(self._bufs) > 1:
            self._cur ^= 1
            if self._inflight[self._cur]:          # the other buffer is still being written: wait before reuse
                self._wait_all()
        else:
            self._wait_all()

    def _finish(self):
        if getattr(self, "_closed", False):
            return
        if self._fill:
            self._submit_current(full=False)
        self._wait_all()
        self._closed = True
        if self._stats[AIO_WRITE_SEC_KEY] > 0:
            self._stats[AIO_SPEED_KEY] = self._stats[AIO_WRITE_BYTES_KEY] / self._stats[AIO_WRITE_SEC_KEY] / 1024 ** 3
