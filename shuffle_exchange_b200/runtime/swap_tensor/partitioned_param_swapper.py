"""Parameter shards on NVMe (ZeRO-Infinity; reference: ``runtime/swap_tensor/partitioned_param_swapper.py:37-434``).

Every parameter shard (``param.ds_tensor`` in the reference, the per-unit low-precision shard here) has a swap file;
``swap_in`` reads shards into slots of one big pinned buffer and marks them ``AVAILABLE`` (optionally asynchronously
with ``synchronize_reads``), ``swap_out_and_release`` writes them back and recycles the slots.
"""
from __future__ import annotations

import os
from enum import Enum
from typing import Dict, List

import torch

from ... import comm as dist
from ...ops.aio import aio_handle
from ...utils.logging import logger
from .constants import *  # noqa: F401,F403
from .utils import MIN_AIO_BYTES, AIO_ALIGNED_BYTES, swap_in_tensors, swap_out_tensors


class PartitionedParamStatus(Enum):
    AVAILABLE = 1          # shard is in (pinned) memory
    NOT_AVAILABLE = 2      # shard is only on NVMe
    INFLIGHT = 3           # read submitted, not yet waited on


class AsyncPartitionedParameterSwapper:
    def __init__(self, ds_config, model_dtype):
        aio = ds_config.aio_config if hasattr(ds_config, "aio_config") else ds_config["aio"]
        param_cfg = ds_config.zero_config.offload_param if hasattr(ds_config, "zero_config") else ds_config["offload_param"]
        self.dtype = model_dtype
        self.aio_config = aio
        rank = dist.get_rank() if dist.is_initialized() else 0
        nvme = str(getattr(param_cfg, "nvme_path", None) or param_cfg["nvme_path"])
        self.swap_folder = os.path.join(nvme, "zero_stage_3", f"{str(model_dtype).split('.')[-1]}params", f"rank{rank}")
        os.makedirs(self.swap_folder, exist_ok=True)
        self.swap_element_size = torch.tensor([], dtype=model_dtype).element_size()
        self.min_aio_bytes = max(MIN_AIO_BYTES, aio[AIO_BLOCK_SIZE])
        self.aligned_bytes = AIO_ALIGNED_BYTES * aio[AIO_INTRA_OP_PARALLELISM]
        self.numel_alignment = self.aligned_bytes // self.swap_element_size
        buffer_size = int(getattr(param_cfg, "buffer_size", None) or param_cfg.get("buffer_size", 1 << 27))
        self.elements_per_buffer = self._io_aligned_numel(buffer_size)
        self.param_buffer_count = int(getattr(param_cfg, "buffer_count", None) or param_cfg.get("buffer_count", 5))
        pin = torch.cuda.is_available()
        self.buffers = torch.empty(self.elements_per_buffer * self.param_buffer_count, dtype=model_dtype, pin_memory=pin)
        self.available_buffer_ids = list(range(self.param_buffer_count))
        self.reserved_buffer_ids: List[int] = []
        self.id_to_path: Dict[int, str] = {}
        self.param_id_to_buffer_id: Dict[int, int] = {}
        self.param_id_to_swap_buffer: Dict[int, torch.Tensor] = {}
        self.param_id_to_numel: Dict[int, int] = {}
        self.pending_writes = 0
        self.pending_reads = 0
        self.inflight_params: List = []
        self.inflight_swap_in_buffers: List[torch.Tensor] = []
        self.available_params = set()
        self.available_numel = 0
        self.aio_read_handle = aio_handle(aio[AIO_BLOCK_SIZE], aio[AIO_QUEUE_DEPTH], aio[AIO_SINGLE_SUBMIT],
                                          aio[AIO_OVERLAP_EVENTS], aio[AIO_INTRA_OP_PARALLELISM])
        self.aio_write_handle = aio_handle(aio[AIO_BLOCK_SIZE], aio[AIO_QUEUE_DEPTH], aio[AIO_SINGLE_SUBMIT],
                                           aio[AIO_OVERLAP_EVENTS], aio[AIO_INTRA_OP_PARALLELISM])
        self.invalid_buffer = torch.tensor(1).to(model_dtype)
        if rank == 0:
            logger.info(f"param swapper: {self.param_buffer_count} buffers x {self.elements_per_buffer} elements at {self.swap_folder}")

    # --------------------------------------------------------------------------------- helpers
    def _io_aligned_numel(self, numel):
        r = numel % self.numel_alignment
        return numel if r == 0 else numel + self.numel_alignment - r

    def swappable_tensor(self, param=None, numel=None):
        n = numel if numel is not None else self._shard(param).numel()
        return self.min_aio_bytes <= n * self.swap_element_size

    def get_path(self, param, must_exist=False):
        return self._get_swap_paths([param], must_exist=must_exist)[0]

    @staticmethod
    def _pid(param):
        return getattr(param, "ds_id", id(param))

    @staticmethod
    def _shard(param):
        return param.ds_tensor if hasattr(param, "ds_tensor") else param

    def _get_swap_paths(self, params, must_exist=False):
        out = []
        for p in params:
            pid = self._pid(p)
            path = self.id_to_path.get(pid)
            if path is None:
                assert not must_exist, f"Path for param id {pid} does not exist"
                path = self.id_to_path[pid] = os.path.join(self.swap_folder, f"{pid}_param.tensor.swp")
            out.append(path)
        return out

    def available_swap_in_buffers(self):
        return len(self.available_buffer_ids)

    def _allocate_and_return_buffers_for_swap_in(self, params):
        compute, swap = [], []
        for p in params:
            pid = self._pid(p)
            assert pid in self.param_id_to_numel, f"param {pid} was never swapped out"
            assert pid not in self.param_id_to_buffer_id, f"param {pid} already assigned swap buffer id {self.param_id_to_buffer_id[pid]}"
            assert self.available_buffer_ids, "no free parameter swap buffers"
            bid = self.available_buffer_ids.pop()
            self.param_id_to_buffer_id[pid] = bid
            n = self.param_id_to_numel[pid]
            aligned = self._io_aligned_numel(n)
            sb = self.buffers.narrow(0, bid * self.elements_per_buffer, aligned)
            self.param_id_to_swap_buffer[pid] = sb
            compute.append(sb.narrow(0, 0, n))
            swap.append(sb)
        return compute, swap

    # ------------------------------------------------------------------------------------- API
    def synchronize_writes(self):
        if self.pending_writes == 0:
            return
        assert self.pending_writes == self.aio_write_handle.wait()
        self.pending_writes = 0
        self.remove_partition_and_release_buffers(self._writing)
        self._writing = []

    def synchronize_reads(self):
        if self.pending_reads == 0:
            return
        assert self.pending_reads == self.aio_read_handle.wait()
        self.pending_reads = 0
        for p, sb in zip(self.inflight_params, self.inflight_swap_in_buffers):
            n = self.param_id_to_numel[self._pid(p)]
            self._shard(p).data = sb.data.narrow(0, 0, n)
            self._shard(p).status = PartitionedParamStatus.AVAILABLE
        self.available_params.update(self._pid(p) for p in self.inflight_params)
        self.available_numel += sum(self.param_id_to_numel[self._pid(p)] for p in self.inflight_params)
        self.inflight_params, self.inflight_swap_in_buffers = [], []

    def remove_partition_and_release_buffers(self, params):
        for p in params:
            pid = self._pid(p)
            bid = self.param_id_to_buffer_id.pop(pid, None)
            if bid is not None:
                self.available_buffer_ids.append(bid)
                self.param_id_to_swap_buffer.pop(pid, None)
                if pid in self.available_params:
                    self.available_params.remove(pid)
                    self.available_numel -= self.param_id_to_numel[pid]
            self._shard(p).data = self.invalid_buffer.data
            self._shard(p).status = PartitionedParamStatus.NOT_AVAILABLE

    def _swap_out(self, params, async_op=False):
        paths = self._get_swap_paths(params)
        bufs = []
        for p in params:
            pid = self._pid(p)
            shard = self._shard(p)
            n = shard.numel()
            self.param_id_to_numel[pid] = n
            sb = self.param_id_to_swap_buffer.get(pid)
            if sb is None:                     # first swap-out: stage through a free slot
                assert self.available_buffer_ids, "no free parameter swap buffers"
                bid = self.available_buffer_ids.pop()
                self.param_id_to_buffer_id[pid] = bid
                sb = self.buffers.narrow(0, bid * self.elements_per_buffer, self._io_aligned_numel(n))
                self.param_id_to_swap_buffer[pid] = sb
                sb.narrow(0, 0, n).copy_(shard.data.reshape(-1))
            bufs.append(sb)
        swap_out_tensors(self.aio_write_handle, bufs, paths)
        self.pending_writes += len(bufs)
        self._writing = list(getattr(self, "_writing", [])) + list(params)
        if not async_op:
            self.synchronize_writes()

    def swap_out_and_release(self, params, async_op=False, force_buffer_release=False):
        if async_op:
            assert force_buffer_release, "Should not release preallocated buffers without completing the swap out. Set force_buffer_release to True to do it anyways"
        self._swap_out(params, async_op=async_op)

    def swap_in(self, params, async_op=True, swap_in_buffers=None):
        assert all(getattr(self._shard(p), "status", PartitionedParamStatus.NOT_AVAILABLE) == PartitionedParamStatus.NOT_AVAILABLE
                   for p in params), "Some params are already available or in flight"
        paths = self._get_swap_paths(params, must_exist=True)
        if swap_in_buffers is None:
            if len(self.available_buffer_ids) < len(params):
                logger.warning(f"Not enough swap in buffers {len(self.available_buffer_ids)} for {len(params)} params")
                return False
            compute, swap = self._allocate_and_return_buffers_for_swap_in(params)
        else:
            swap = [b.narrow(0, 0, self._io_aligned_numel(self.param_id_to_numel[self._pid(p)])) for b, p in zip(swap_in_buffers, params)]
        swap_in_tensors(self.aio_read_handle, swap, paths)
        self.pending_reads += len(params)
        self.inflight_params += list(params)
        self.inflight_swap_in_buffers += swap
        for p in params:
            self._shard(p).status = PartitionedParamStatus.INFLIGHT
        if not async_op:
            self.synchronize_reads()
        return True

    def swap_into_buffer(self, param, dest_buffer):
        pid = self._pid(param)
        n = self.param_id_to_numel[pid]
        require = self._io_aligned_numel(n)
        assert dest_buffer.numel() >= require and dest_buffer.device.type == "cpu"
        swap_in_tensors(self.aio_read_handle, [dest_buffer.narrow(0, 0, require)], [self.get_path(param, True)])
        assert self.aio_read_handle.wait() == 1
        self._shard(param).data = dest_buffer.narrow(0, 0, n).data
        self._shard(param).status = PartitionedParamStatus.AVAILABLE

    def get_buffer(self, param, numel):
        pid = self._pid(param)
        assert self.available_buffer_ids, f"No swap buffers to allocate for fp16 param {pid}"
        bid = self.available_buffer_ids.pop()
        self.param_id_to_buffer_id[pid] = bid
        self.param_id_to_numel[pid] = numel
        sb = self.buffers.narrow(0, bid * self.elements_per_buffer, self._io_aligned_numel(numel))
        self.param_id_to_swap_buffer[pid] = sb
        return sb.narrow(0, 0, numel)

    def reserve_available_buffers(self):
        out = []
        for bid in self.available_buffer_ids:
            out.append(self.buffers.narrow(0, bid * self.elements_per_buffer, self.elements_per_buffer))
            self.reserved_buffer_ids.append(bid)
        self.available_buffer_ids = []
        return out

    def release_reserved_buffers(self):
        self.available_buffer_ids += self.reserved_buffer_ids
        self.reserved_buffer_ids = []
