"""CUDA-only device helper.

The reference routes every device call through a multi-vendor ``get_accelerator()`` object
(SURVEY.md §2.3 N22 lists the ~60 methods used).  This framework targets B200 only, so the same
surface is a thin module-level object over ``torch.cuda``; on a machine without a GPU (unit tests
under gloo) it degrades to CPU so host-side logic stays testable.
"""
from __future__ import annotations

import contextlib
import os

import torch


class _Device:
    _name = "cuda"

    # ---- identity -------------------------------------------------------------------
    def is_available(self) -> bool:
        return torch.cuda.is_available()

    def device_name(self, index=None) -> str:
        if not self.is_available():
            return "cpu"
        return "cuda" if index is None else f"cuda:{index}"

    def current_device(self) -> int:
        return torch.cuda.current_device() if self.is_available() else 0

    def current_device_name(self) -> str:
        return f"cuda:{torch.cuda.current_device()}" if self.is_available() else "cpu"

    def device(self, index=None):
        return torch.device(self.device_name(index))

    def set_device(self, index):
        if self.is_available():
            torch.cuda.set_device(index)

    def device_count(self) -> int:
        return torch.cuda.device_count() if self.is_available() else 0

    def communication_backend_name(self) -> str:
        return "nccl" if self.is_available() else "gloo"

    def is_synchronized_device(self) -> bool:
        return not self.is_available()

    def resolves_data_dependency(self) -> bool:
        return not self.is_available()

    def handles_memory_backpressure(self) -> bool:
        return not self.is_available()

    def on_accelerator(self, tensor) -> bool:
        return tensor.is_cuda

    # ---- streams / events -----------------------------------------------------------
    def synchronize(self, index=None):
        if self.is_available():
            torch.cuda.synchronize(index)

    def Stream(self, **kw):
        return torch.cuda.Stream(**kw) if self.is_available() else None

    def Event(self, **kw):
        return torch.cuda.Event(**kw) if self.is_available() else None

    def current_stream(self, index=None):
        return torch.cuda.current_stream(index) if self.is_available() else None

    def default_stream(self, index=None):
        return torch.cuda.default_stream(index) if self.is_available() else None

    def stream(self, s):
        if s is None or not self.is_available():
            return contextlib.nullcontext()
        return torch.cuda.stream(s)

    # ---- memory -----------------------------------------------------------------------
    def empty_cache(self):
        if self.is_available():
            torch.cuda.empty_cache()

    def memory_allocated(self, index=None) -> int:
        return torch.cuda.memory_allocated(index) if self.is_available() else 0

    def max_memory_allocated(self, index=None) -> int:
        return torch.cuda.max_memory_allocated(index) if self.is_available() else 0

    def memory_reserved(self, index=None) -> int:
        return torch.cuda.memory_reserved(index) if self.is_available() else 0

    def max_memory_reserved(self, index=None) -> int:
        return torch.cuda.max_memory_reserved(index) if self.is_available() else 0

    def reset_peak_memory_stats(self, index=None):
        if self.is_available():
            torch.cuda.reset_peak_memory_stats(index)

    def total_memory(self, index=None) -> int:
        if not self.is_available():
            import psutil
            return psutil.virtual_memory().total
        return torch.cuda.get_device_properties(index or 0).total_memory

    def available_memory(self, index=None) -> int:
        if not self.is_available():
            import psutil
            return psutil.virtual_memory().available
        free, _ = torch.cuda.mem_get_info(index)
        return free

    def pin_memory(self, tensor, align_bytes=1):
        return tensor.pin_memory() if self.is_available() else tensor

    def is_pinned(self, tensor) -> bool:
        return tensor.is_pinned()

    # ---- RNG ----------------------------------------------------------------------------
    def manual_seed(self, seed):
        torch.manual_seed(seed)
        if self.is_available():
            torch.cuda.manual_seed(seed)

    def manual_seed_all(self, seed):
        torch.manual_seed(seed)
        if self.is_available():
            torch.cuda.manual_seed_all(seed)

    def get_rng_state(self, index=None):
        return torch.cuda.get_rng_state(index or "cuda") if self.is_available() else torch.get_rng_state()

    def set_rng_state(self, state, index=None):
        if self.is_available():
            torch.cuda.set_rng_state(state, index or "cuda")
        else:
            torch.set_rng_state(state)

    # ---- profiling ranges -----------------------------------------------------------------
    def range_push(self, msg):
        if self.is_available():
            torch.cuda.nvtx.range_push(msg)

    def range_pop(self):
        if self.is_available():
            torch.cuda.nvtx.range_pop()

    # ---- graphs -----------------------------------------------------------------------------
    def create_graph(self):
        return torch.cuda.CUDAGraph()

    def capture_to_graph(self, graph, pool=None, stream=None):
        return torch.cuda.graph(graph, pool=pool, stream=stream)

    def replay_graph(self, graph):
        graph.replay()

    # ---- dtype support ---------------------------------------------------------------------
    def is_bf16_supported(self) -> bool:
        return True

    def is_fp16_supported(self) -> bool:
        return True

    def supported_dtypes(self):
        return [torch.float32, torch.float16, torch.bfloat16]

    def lazy_call(self, cb):
        cb()


_DEVICE = _Device()


def get_accelerator() -> _Device:
    """Name kept for drop-in compatibility with scripts written against the reference."""
    return _DEVICE


def local_device() -> torch.device:
    if _DEVICE.is_available():
        return torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")) % max(1, torch.cuda.device_count()))
    return torch.device("cpu")
