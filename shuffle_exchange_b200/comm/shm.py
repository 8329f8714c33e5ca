"""Shared-memory all-reduce for CPU tensors among the ranks of one host (C++: ``csrc/host/shm_comm.cpp``).

Reference: ``ShareMemCommBuilder`` / ``torch.ops.deepspeed.inference_all_reduce_`` (``comm/torch.py:29-35,166-171``): CPU
inference with tensor parallelism spends its communication time in small all-reduces for which gloo's TCP loop-back is
an order of magnitude slower than a memcpy through ``/dev/shm``.  ``inference_all_reduce`` in ``comm.comm`` uses this
path when every rank of the (world) group lives on this host; anything else falls back to ``all_reduce``."""
from __future__ import annotations

import ctypes
import os
from typing import Optional

import torch

from ..native import host as _host

_DT = {torch.float32: 0, torch.bfloat16: 1, torch.float16: 2}
_host.register({
    "sxbh_shm_init": [ctypes.c_char_p, ctypes.c_int, ctypes.c_int, ctypes.c_longlong],
    "sxbh_shm_allreduce": [ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int],
    "sxbh_shm_barrier": [],
    "sxbh_shm_finalize": [],
})


class ShmAllReduce:
    """One instance per process; ``initialize`` is collective over the ranks of the host."""

    def __init__(self):
        self.initialized = False
        self.world = self.rank = None

    def initialize(self, world: int, rank: int, name: Optional[str] = None, slot_bytes: int = 32 << 20) -> bool:
        if self.initialized:
            return True
        name = name or f"sxb200_shm_{os.environ.get('MASTER_PORT', '0')}_{os.getuid()}"
        rc = _host.load().sxbh_shm_init(name.encode(), int(world), int(rank), int(slot_bytes))
        if rc != 0:
            raise OSError(-rc, f"shared-memory segment {name}: {os.strerror(-rc)}")
        self.initialized, self.world, self.rank = True, world, rank
        return True

    @staticmethod
    def supports(tensor: torch.Tensor) -> bool:
        return tensor.device.type == "cpu" and tensor.dtype in _DT and tensor.is_contiguous()

    def all_reduce(self, tensor: torch.Tensor) -> torch.Tensor:
        """In-place sum (fp32 accumulation for 16-bit types)."""
        if not self.initialized:
            raise RuntimeError("ShmAllReduce.initialize was not called")
        if not self.supports(tensor):
            raise TypeError(f"shared-memory all-reduce needs a contiguous CPU fp32/bf16/fp16 tensor, got "
                            f"{tensor.dtype} on {tensor.device}")
        rc = _host.load().sxbh_shm_allreduce(ctypes.c_void_p(tensor.data_ptr()), tensor.numel(), _DT[tensor.dtype])
        if rc != 0:
            raise OSError(-rc, os.strerror(-rc))
        return tensor

    def barrier(self) -> None:
        _host.load().sxbh_shm_barrier()

    def finalize(self) -> None:
        if self.initialized:
            _host.load().sxbh_shm_finalize()
            self.initialized = False


_INSTANCE: Optional[ShmAllReduce] = None


def get_shm_comm() -> ShmAllReduce:
    global _INSTANCE
    if _INSTANCE is None:
        _INSTANCE = ShmAllReduce()
    return _INSTANCE


def all_ranks_on_this_host() -> bool:
    """True when LOCAL_SIZE / LOCAL_WORLD_SIZE says the whole job runs here (the launcher exports them)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_WORLD_SIZE", os.environ.get("LOCAL_SIZE", "0")) or 0)
    return world > 1 and local == world
