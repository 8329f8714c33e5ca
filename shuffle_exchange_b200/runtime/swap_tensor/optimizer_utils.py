"""NVMe residency of optimizer state.

Reference: ``runtime/swap_tensor/optimizer_utils.py:20-475`` keeps, per *parameter*, a ``OptimizerStateSwapInfo``
(list of state tensors, their swap files, unpinned gradient fragments).  In this framework optimizer state is held per
FLAT GROUP as whole fp32 shards (master, exp_avg, exp_avg_sq, ...), so the unit of swapping is one flat group: a handful
of large sequential files instead of thousands of small ones — the access pattern NVMe is fastest at.
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional

import torch

from ... import comm as dist
from ...ops.aio import aio_handle
from ...utils.logging import logger
from .constants import *  # noqa: F401,F403
from .utils import MIN_AIO_BYTES, SwapBufferManager, swap_in_tensors, swap_out_tensors


class FlattenedTensorSwapInfo:
    def __init__(self, path, length, offset):
        self.path, self.length, self.offset = path, length, offset


class OptimizerStateSwapInfo:
    """Swap bookkeeping of one flat group (``key`` = the group's fp32 master parameter)."""

    def __init__(self, parameter: torch.Tensor, numel: int, base_folder: str, key_id: int):
        self.tensors: Dict[str, Optional[torch.Tensor]] = {}
        self.param_id = key_id
        self.swap_folder = base_folder
        self.swap_paths: Dict[str, str] = {}
        self.swapped_gradients: Dict[int, FlattenedTensorSwapInfo] = {}
        self.unswapped_gradients: Dict[int, torch.Tensor] = {}
        self.tensor_numel = numel
        self.tensor_dtype = parameter.dtype
        self.tensor_device = parameter.device
        self.has_state_tensors = False

    def numel(self):
        return self.tensor_numel

    def has_gradients(self):
        return bool(self.swapped_gradients) or bool(self.unswapped_gradients)

    def path_of(self, name: str) -> str:
        p = self.swap_paths.get(name)
        if p is None:
            p = self.swap_paths[name] = os.path.join(self.swap_folder, f"{self.param_id}_{name}.tensor.swp")
        return p

    def get_swap_paths(self):
        return list(self.swap_paths.values())


class OptimizerSwapper:
    """Common part of the optimizer swappers: aio handle, aligned element counts, pinned buffers, file layout
    ``<nvme_path>/zero_stage_3/optimizer/rank<r>/``."""

    def __init__(self, swap_config, aio_config, base_folder, optimizer, largest_numel, device, dtype, timers):
        self.swap_config = swap_config
        self.aio_config = aio_config
        self.swap_element_size = torch.tensor([], dtype=dtype).element_size()
        rank = dist.get_rank() if dist.is_initialized() else 0
        self.swap_folder = os.path.join(base_folder, "optimizer", f"rank{rank}")
        os.makedirs(self.swap_folder, exist_ok=True)
        self.optimizer = optimizer
        self.min_aio_bytes = max(MIN_AIO_BYTES, aio_config[AIO_BLOCK_SIZE])
        from .utils import AIO_ALIGNED_BYTES
        self.aligned_bytes = AIO_ALIGNED_BYTES * aio_config[AIO_INTRA_OP_PARALLELISM]
        self.numel_alignment = self.aligned_bytes // self.swap_element_size
        self.largest_numel = self._io_aligned_numel(largest_numel)
        self.dtype = dtype
        buffer_count = getattr(swap_config, "buffer_count", 4)
        self.swap_buffer_manager = SwapBufferManager(num_elems=self.largest_numel, count=buffer_count, dtype=dtype)
        self.swap_params_info: Dict[int, OptimizerStateSwapInfo] = {}
        self.timers = timers
        self.timer_names = set()
        self.aio_handle = aio_handle(aio_config[AIO_BLOCK_SIZE], aio_config[AIO_QUEUE_DEPTH], aio_config[AIO_SINGLE_SUBMIT],
                                     aio_config[AIO_OVERLAP_EVENTS], aio_config[AIO_INTRA_OP_PARALLELISM])

    def _io_aligned_numel(self, numel):
        r = numel % self.numel_alignment
        return numel if r == 0 else numel + self.numel_alignment - r

    def swappable_tensor(self, param=None, numel=None):
        assert param is not None or numel is not None
        n = param.numel() if param is not None else numel
        return self.min_aio_bytes <= n * self.swap_element_size

    def _info(self, key: torch.Tensor, create: bool = True) -> Optional[OptimizerStateSwapInfo]:
        kid = id(key)
        info = self.swap_params_info.get(kid)
        if info is None and create:
            info = self.swap_params_info[kid] = OptimizerStateSwapInfo(key, key.numel(), self.swap_folder, len(self.swap_params_info))
        return info

    def purge_state(self):
        for info in self.swap_params_info.values():
            info.tensors = {}

    def get_swap_paths(self, key):
        info = self._info(key, create=False)
        return info.get_swap_paths() if info else []
