"""Synchronous optimizer-state swapper (reference: ``runtime/swap_tensor/partitioned_optimizer_swapper.py:28-228``):
``swap_in_optimizer_state(key)`` brings the fp32 master and all optimizer moments of one flat group from NVMe into
pinned buffers (and re-points the live tensors at them); ``swap_out_optimizer_state(key)`` writes them back and frees
the buffers.  Gradients that arrive while the state is on disk are written next to it by ``swap_out_gradients``."""
from __future__ import annotations

from typing import Dict, List

import torch

from ...utils.logging import logger
from .async_swapper import AsyncTensorSwapper
from .optimizer_utils import OptimizerSwapper
from .utils import get_sized_buffers, swap_in_tensors, swap_out_tensors

DEBUG_MODE = False
SWAP_IN_PARAM_TIMER, SWAP_OUT_PARAM_TIMER = "swap_in_param", "swap_out_param"
SWAP_IN_GRADIENT_TIMER = "swap_in_gradient"


class PartitionedOptimizerSwapper(OptimizerSwapper):
    def __init__(self, swap_config, aio_config, base_folder, optimizer, largest_numel, device, dtype, timers):
        super().__init__(swap_config, aio_config, base_folder, optimizer, largest_numel, device, dtype, timers)
        self.gradient_swapper = AsyncTensorSwapper(aio_handle=self.aio_handle, numel_alignment=self.numel_alignment,
                                                   timers=timers)
        self._resident: Dict[int, List[torch.Tensor]] = {}

    def _state_tensors(self, key: torch.Tensor) -> Dict[str, torch.Tensor]:
        """name -> tensor for the master itself and every tensor optimizer state of ``key``."""
        out = {"fp32": key}
        for name, v in self.optimizer.state.get(key, {}).items():
            if torch.is_tensor(v) and v.numel() == key.numel():
                out[name] = v
        return out

    def initialize_parameters(self, parameters, src_tensors):
        for p, src in zip(parameters, src_tensors):
            p.data.copy_(src.data)

    def initialize_from_swapped_fp16_params(self, fp16_partitions_info, fp16_num_elems, fp16_pinned_buffers, fp32_parameters):
        raise NotImplementedError("low-precision shards of this framework never live on NVMe without their masters")

    def flush_gradients(self):
        self.gradient_swapper._flush_buffers_until_complete() if self.gradient_swapper.has_buffers() else None

    # ---------------------------------------------------------------------------------- state
    def swap_out_optimizer_state(self, parameter, async_swap=False):
        info = self._info(parameter)
        tensors = self._state_tensors(parameter)
        n = parameter.numel()
        aligned = self._io_aligned_numel(n)
        bufs = self._resident.pop(id(parameter), None)
        own = bufs is None
        if own:
            bufs = self.swap_buffer_manager.allocate(num_elems=aligned, count=len(tensors), dtype=parameter.dtype)
            assert bufs is not None, "No swap buffers to allocate optimizer state"
            for b, t in zip(bufs, tensors.values()):
                b.narrow(0, 0, n).copy_(t.reshape(-1))
        paths = [info.path_of(name) for name in tensors]
        swap_out_tensors(self.aio_handle, bufs, paths)
        assert self.aio_handle.wait() == len(bufs)
        self.swap_buffer_manager.free(bufs)
        info.has_state_tensors = True
        info.tensors = {name: None for name in tensors}
        # release device/host memory of the live tensors: they now alias an empty storage until swapped in
        for name, t in tensors.items():
            t.data = torch.empty(0, dtype=t.dtype, device=t.device)

    def swap_in_optimizer_state(self, parameter, async_parameter=None):
        info = self._info(parameter, create=False)
        assert info is not None and info.has_state_tensors, "state was never swapped out"
        names = list(info.tensors.keys())
        n = info.numel()
        aligned = self._io_aligned_numel(n)
        bufs = self.swap_buffer_manager.allocate(num_elems=aligned, count=len(names), dtype=info.tensor_dtype)
        assert bufs is not None, f"PartitionedOptimizerSwapper ran out of swap buffers (need {len(names)})"
        swap_in_tensors(self.aio_handle, bufs, [info.path_of(nm) for nm in names])
        assert self.aio_handle.wait() == len(bufs)
        live = {"fp32": parameter}
        live.update({k: v for k, v in self.optimizer.state.get(parameter, {}).items() if torch.is_tensor(v)})
        for nm, b in zip(names, bufs):
            live[nm].data = b.narrow(0, 0, n)
        self._resident[id(parameter)] = bufs

    # ------------------------------------------------------------------------------ gradients
    def swap_out_gradients(self, parameter, gradient_offsets, gradient_tensors):
        info = self._info(parameter)
        if not self.gradient_swapper.has_buffers():
            pinned = self.swap_buffer_manager.allocate(self.largest_numel, 2, self.dtype)
            assert pinned is not None
            self.gradient_swapper.add_buffers(pinned)
        from .optimizer_utils import FlattenedTensorSwapInfo
        import os
        paths = []
        for off, g in zip(gradient_offsets, gradient_tensors):
            if not self.swappable_tensor(g):
                info.unswapped_gradients[off] = g
                continue
            path = os.path.join(self.swap_folder, f"{info.param_id}_gradient_{off}_{g.numel()}.tensor.swp")
            info.swapped_gradients[off] = FlattenedTensorSwapInfo(path, g.numel(), off)
            paths.append((g, path))
        if paths:
            self.gradient_swapper.swap_out_tensors([g for g, _ in paths], [p for _, p in paths])

    def swap_in_gradients(self, parameter, dest_buffer):
        info = self._info(parameter, create=False)
        if info is None or not info.has_gradients():
            return
        self.flush_gradients()
        for off, sw in info.swapped_gradients.items():
            aligned = self._io_aligned_numel(sw.length)
            buf = self.swap_buffer_manager.allocate(aligned, 1, dest_buffer.dtype)
            assert buf is not None
            swap_in_tensors(self.aio_handle, buf, [sw.path])
            assert self.aio_handle.wait() == 1
            dest_buffer.narrow(0, off, sw.length).copy_(buf[0].narrow(0, 0, sw.length))
            self.swap_buffer_manager.free(buf)
        for off, g in info.unswapped_gradients.items():
            dest_buffer.narrow(0, off, g.numel()).copy_(g.reshape(-1))
        info.swapped_gradients, info.unswapped_gradients = {}, {}

    def release_swap_buffers(self, parameter):
        bufs = self._resident.pop(id(parameter), None)
        if bufs:
            self.swap_buffer_manager.free(bufs)
