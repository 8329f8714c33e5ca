"""Shared machinery of the 1-bit optimizer family: communication backend selection, per-parameter
error-feedback buffers sized for the compressed all-reduce, and the warm-up -> compression switch that also
turns off the engine's own gradient all-reduce (reference: ``runtime/fp16/onebit/adam.py:70-112, 281-306``)."""
from __future__ import annotations

import numpy as np
import torch

from .... import comm as dist


class OnebitBase(torch.optim.Optimizer):
    def _setup_comm(self, deepspeed, comm_backend_name: str, cuda_aware: bool = False):
        self.deepspeed = deepspeed
        self.comm_backend_name = comm_backend_name
        self.using_pipeline = deepspeed is not None and hasattr(deepspeed, "pipeline_enable_backward_allreduce")
        mpu = getattr(deepspeed, "mpu", None)
        if comm_backend_name in ("nccl", "compressed", "hccl", "mpi"):
            if comm_backend_name in ("hccl", "mpi"):
                raise NotImplementedError(f"comm backend {comm_backend_name} is not available on this platform; use 'nccl'")
            from ...comm.nccl import NcclBackend
            from ...comm.compressed import CompressedBackend
            cls = NcclBackend if comm_backend_name == "nccl" else CompressedBackend
            self.comm_backend_handle = cls(mpu) if dist.is_initialized() else None
        else:
            raise ValueError(f"unknown comm backend {comm_backend_name}")
        self.size = self.comm_backend_handle.size if self.comm_backend_handle is not None else 1
        self.divider = int(self.size * 8 / np.gcd(self.size, 8))

    def _error_buffers(self, state, p):
        """Allocate (once) the worker/server error-feedback buffers, padded so that the flat momentum splits evenly
        into ``world`` byte-aligned sign chunks."""
        if "worker_error" in state:
            return
        n = p.numel()
        unit = self.size * self.divider
        padded = n + ((-n) % unit)
        state["tensor_size"], state["corrected_tensor_size"] = n, padded
        state["server_chunk_size"] = padded // self.size
        state["worker_error"] = torch.zeros(padded, device=p.device, dtype=p.dtype)
        state["server_error"] = torch.zeros(padded // self.size, device=p.device, dtype=p.dtype)

    def _compressed_allreduce(self, tensor, state):
        if self.size <= 1 or self.comm_backend_handle is None:
            return tensor
        local_rank = getattr(self.deepspeed, "local_rank", 0)
        return self.comm_backend_handle.compressed_allreduce(tensor, state["worker_error"], state["server_error"],
                                                             local_rank)

    def _set_engine_allreduce(self, enabled: bool):
        if self.deepspeed is None:
            return
        if self.using_pipeline:
            self.deepspeed.pipeline_enable_backward_allreduce = enabled
        else:
            self.deepspeed.enable_backward_allreduce = enabled

    @staticmethod
    def _drop_error_buffers(state):
        state.pop("worker_error", None)
        state.pop("server_error", None)
