"""Low-precision parameter shards on NVMe (ZeRO-Infinity ``offload_param.device = nvme``).

Reference: ``runtime/swap_tensor/partitioned_param_swapper.py`` — every rank's ``ds_tensor`` partitions live in swap
files and pass through a fixed pool of pinned buffers.  Here the unit of swapping is the unit of everything else in
this engine: the flat group's shard (one file per group and rank).  :class:`NvmeShardStore` keeps ``buffer_count``
shards resident in (pinned) host memory — more while the resident elements stay within ``max_in_cpu``
(``offload_param.max_in_cpu``: how much of the model may stay in host memory) — reads a shard with the asynchronous file engine
(``csrc/host/aio.cpp``) when someone asks for it and writes it back on eviction if it was modified since it was read.
The flat group sees an ordinary host tensor (``FlatGroup.lp_shard`` becomes a property over the store), so gathers,
optimizer write-back and checkpointing need no NVMe-specific code; they only say when they have written
(:meth:`FlatGroup.lp_shard_written`).
"""
from __future__ import annotations

import os
from collections import OrderedDict
from typing import Dict, Tuple

import torch

from ...ops.aio import aio_handle

_IO_ALIGN = 4096          # bytes: files are padded to whole blocks so O_DIRECT transfers stay aligned


class NvmeShardStore:
    def __init__(self, folder: str, dtype: torch.dtype, buffer_count: int = 5, aio_config: dict = None,
                 pin_memory: bool = False, max_in_cpu: int = 0):
        self.folder = folder
        os.makedirs(folder, exist_ok=True)
        self.dtype = dtype
        self.capacity = max(1, int(buffer_count))
        self.max_in_cpu = max(0, int(max_in_cpu))       # elements that may be resident beyond the first `capacity` shards
        self.pin = bool(pin_memory) and torch.cuda.is_available()
        a = aio_config or {}
        self.handle = aio_handle(int(a.get("block_size", 1 << 20)), int(a.get("queue_depth", 8)),
                                 bool(a.get("single_submit", False)), bool(a.get("overlap_events", True)),
                                 int(a.get("intra_op_parallelism", a.get("thread_count", 1))))
        self._esize = torch.tensor([], dtype=dtype).element_size()
        self._numel: Dict[str, int] = {}
        self._resident: "OrderedDict[str, Tuple[torch.Tensor, bool]]" = OrderedDict()      # key -> (padded buffer, dirty)
        self.reads = self.writes = 0

    # ------------------------------------------------------------------------------------ layout
    def _padded(self, numel: int) -> int:
        per = _IO_ALIGN // self._esize
        return (numel + per - 1) // per * per

    def path(self, key: str) -> str:
        return os.path.join(self.folder, f"{key}.swp")

    def _buffer(self, numel: int) -> torch.Tensor:
        t = torch.zeros(self._padded(numel), dtype=self.dtype)
        return t.pin_memory() if self.pin else t

    # --------------------------------------------------------------------------------------- api
    def put(self, key: str, shard: torch.Tensor) -> None:
        """Create / replace the file of ``key`` with ``shard`` (not kept resident)."""
        self._numel[key] = shard.numel()
        buf = self._buffer(shard.numel())
        buf[: shard.numel()].copy_(shard.detach().reshape(-1).to("cpu"))
        self._write(key, buf)
        self._resident.pop(key, None)

    def get(self, key: str) -> torch.Tensor:
        """The shard as a host tensor (a view of its resident buffer); most recently used last."""
        hit = self._resident.get(key)
        if hit is None:
            while self._resident and not self._room_for(self._numel[key]):
                self._evict_one()
            buf = self._buffer(self._numel[key])
            rc = self.handle.sync_pread(buf, self.path(key))
            if rc < 0:
                raise IOError(f"reading parameter shard {self.path(key)} failed ({rc})")
            self.reads += 1
            hit = (buf, False)
            self._resident[key] = hit
        else:
            self._resident.move_to_end(key)
        return hit[0][: self._numel[key]]

    def written(self, key: str) -> None:
        """The resident copy of ``key`` was modified: it must reach the file before its buffer is reused."""
        if key in self._resident:
            self._resident[key] = (self._resident[key][0], True)
            self._resident.move_to_end(key)

    def flush(self) -> None:
        for key, (buf, dirty) in list(self._resident.items()):
            if dirty:
                self._write(key, buf)
                self._resident[key] = (buf, False)

    def release_all(self) -> None:
        self.flush()
        self._resident.clear()

    @property
    def resident_count(self) -> int:
        return len(self._resident)

    @property
    def resident_numel(self) -> int:
        return sum(self._numel[k] for k in self._resident)

    # ---------------------------------------------------------------------------------- internal
    def _room_for(self, numel: int) -> bool:
        if len(self._resident) < self.capacity:
            return True
        return self.max_in_cpu > 0 and self.resident_numel + numel <= self.max_in_cpu

    def _evict_one(self) -> None:
        key, (buf, dirty) = self._resident.popitem(last=False)
        if dirty:
            self._write(key, buf)

    def _write(self, key: str, buf: torch.Tensor) -> None:
        rc = self.handle.sync_pwrite(buf, self.path(key))
        if rc < 0:
            raise IOError(f"writing parameter shard {self.path(key)} failed ({rc})")
        self.writes += 1
