"""Runtime helpers (reference: deepspeed/runtime/utils.py — ``clip_grad_norm_``,
``get_global_norm_of_tensors``, ``partition_uniform/balanced``, ``see_memory_usage`` :771-793,
``all_gather_dp_groups`` :972-1007, ``CheckOverflow``, ``PartitionedTensor``)."""
from __future__ import annotations

import gc
import math
from bisect import bisect_left
from typing import List

import psutil
import torch

from ..comm import comm as dist
from ..device import get_accelerator
from ..utils import groups
from ..utils.logging import logger


def noop_decorator(func):
    return func


def ensure_directory_exists(filename):
    import os
    os.makedirs(os.path.dirname(filename), exist_ok=True)


def set_random_seed(seed):
    import random
    import numpy
    random.seed(seed)
    numpy.random.seed(seed)
    torch.manual_seed(seed)


def is_model_parallel_parameter(p) -> bool:
    return bool(getattr(p, "model_parallel", False) or getattr(p, "tensor_model_parallel", False))


def get_global_norm(norm_list):
    return math.sqrt(sum(n ** 2.0 for n in norm_list))


def get_global_norm_of_tensors(input_tensors, norm_type=2, mpu=None, use_graph=False, moe_ep_group=None):
    tensors = [t for t in input_tensors if t is not None]
    if not tensors:
        return 0.0
    dev = tensors[0].device
    if norm_type == float("inf"):
        total = torch.stack([t.abs().max().float() for t in tensors]).max()
        if mpu is not None:
            dist.all_reduce(total, op=dist.ReduceOp.MAX, group=mpu.get_model_parallel_group())
        if moe_ep_group is not None:
            dist.all_reduce(total, op=dist.ReduceOp.MAX, group=moe_ep_group)
        return float(total)
    total = torch.zeros(1, device=dev, dtype=torch.float32)
    for t in tensors:
        total += t.float().norm(norm_type) ** norm_type
    if mpu is not None:
        dist.all_reduce(total, op=dist.ReduceOp.SUM, group=mpu.get_model_parallel_group())
    if moe_ep_group is not None:
        dist.all_reduce(total, op=dist.ReduceOp.SUM, group=moe_ep_group)
    v = float(total) ** (1.0 / norm_type)
    if v == float("inf") or v == -float("inf") or v != v:
        return -1.0
    return v


def clip_grad_norm_(parameters, max_norm, norm_type=2, mpu=None):
    """Global-norm clipping that counts model-parallel parameters once (Megatron convention)."""
    if isinstance(parameters, torch.Tensor):
        parameters = [parameters]
    parameters = [p for p in parameters if p.grad is not None]
    if not parameters:
        return 0.0
    grads = []
    for p in parameters:
        if mpu is not None and not is_model_parallel_parameter(p) and groups.get_model_parallel_rank() != 0:
            continue
        grads.append(p.grad.detach())
    total_norm = get_global_norm_of_tensors(grads, norm_type, mpu) if grads else 0.0
    if mpu is not None and not grads:
        total_norm = get_global_norm_of_tensors([torch.zeros(1, device=parameters[0].device)], norm_type, mpu)
    clip_coef = max_norm / (total_norm + 1e-6)
    if clip_coef < 1:
        for p in parameters:
            p.grad.detach().mul_(clip_coef)
    return total_norm


def clip_tensors_by_global_norm(input_tensors, max_norm=1.0, global_norm=None, mpu=None, eps=1e-6, use_graph=False):
    if global_norm is None:
        global_norm = get_global_norm_of_tensors(input_tensors, mpu=mpu)
    clip_coef = max_norm / (global_norm + eps)
    if clip_coef < 1:
        for t in input_tensors:
            t.detach().mul_(clip_coef)
    return global_norm


class CheckOverflow:
    """inf/nan detection over parameter gradients with cross-rank agreement."""

    def __init__(self, param_groups=None, mpu=None, zero_reduce_scatter=False, deepspeed=None):
        self.mpu = mpu
        self.params = [] if param_groups else None
        self.zero_reduce_scatter = zero_reduce_scatter
        self.deepspeed = deepspeed
        self.has_moe_params = False
        if param_groups:
            for g in param_groups:
                for p in g:
                    self.params.append(p)
                    if hasattr(p, "allreduce") and not p.allreduce:
                        self.has_moe_params = True

    @staticmethod
    def _has_inf_or_nan(x):
        s = float(x.float().sum())
        return s in (float("inf"), -float("inf")) or s != s

    def has_overflow_serial(self, params):
        return any(p.grad is not None and self._has_inf_or_nan(p.grad.data) for p in params)

    def check_using_norm(self, norm_group, reduce_overflow=True):
        overflow = -1 in norm_group
        flag = torch.tensor([float(overflow)], device=get_accelerator().current_device_name())
        if reduce_overflow and dist.is_initialized():
            dist.all_reduce(flag, op=dist.ReduceOp.MAX)
        return bool(flag.item())

    def has_overflow(self, params, has_moe_params=None):
        overflow = self.has_overflow_serial(params)
        flag = torch.tensor([float(overflow)], device=get_accelerator().current_device_name())
        if dist.is_initialized():
            if self.mpu is not None:
                dist.all_reduce(flag, op=dist.ReduceOp.MAX, group=self.mpu.get_model_parallel_group())
            dist.all_reduce(flag, op=dist.ReduceOp.MAX)
        return bool(flag.item())

    def check(self, param_groups=None):
        params = [p for g in param_groups for p in g] if param_groups is not None else self.params
        return self.has_overflow(params)


def prefix_sum_inc(weights):
    out = list(weights)
    for i in range(1, len(out)):
        out[i] += out[i - 1]
    return out


def partition_uniform(num_items, num_parts):
    """Boundaries ``parts[p]..parts[p+1]`` splitting ``num_items`` as evenly as possible."""
    parts = [0] * (num_parts + 1)
    if num_items <= num_parts:
        for p in range(num_parts + 1):
            parts[p] = min(p, num_items)
        return parts
    chunk, residual = divmod(num_items, num_parts)
    for p in range(1, num_parts + 1):
        parts[p] = parts[p - 1] + chunk + (1 if p <= residual else 0)
    return parts


def partition_balanced(weights, num_parts):
    """Contiguous partition minimising the heaviest part (binary search on the bottleneck)."""
    n = len(weights)
    if n <= num_parts:
        return partition_uniform(n, num_parts)
    pre = [0] + prefix_sum_inc(weights)

    def feasible(cap):
        parts, start = [0], 0
        for _ in range(num_parts):
            # furthest end such that sum(weights[start:end]) <= cap
            lo = bisect_left(pre, pre[start] + cap + 1e-9, lo=start + 1) - 1
            if lo <= start:
                return None
            start = min(lo, n)
            parts.append(start)
            if start == n:
                break
        if parts[-1] != n:
            return None
        while len(parts) < num_parts + 1:
            parts.append(n)
        return parts

    lo, hi = max(weights), sum(weights)
    best = feasible(hi)
    for _ in range(64):
        if hi - lo < 1e-6 * max(1.0, hi):
            break
        mid = (lo + hi) / 2
        r = feasible(mid)
        if r is not None:
            best, hi = r, mid
        else:
            lo = mid
    return best


class PartitionedTensor:
    """A tensor split (flattened) across a process group with metadata to rebuild it
    (used by the pipeline engine when ``partition_activations`` is on)."""

    def __init__(self, tensor, group, partition_meta=None):
        self.group = group
        self.num_parts = dist.get_world_size(group=group)
        self.rank = dist.get_rank(group=group)
        self.orig_size = list(tensor.size())
        self.orig_device = tensor.device
        self.local_data, self.partition = self._partition_tensor(tensor)

    @classmethod
    def from_meta(cls, meta, local_part, group, device=None):
        assert meta.dtype == torch.long
        obj = cls.__new__(cls)
        obj.group = group
        obj.num_parts = dist.get_world_size(group=group)
        obj.rank = dist.get_rank(group=group)
        m = meta.tolist()
        ndim = m[0]
        obj.orig_size = m[1:1 + ndim]
        obj.partition = m[2 + ndim:]
        obj.orig_device = device or local_part.device
        obj.local_data = local_part
        return obj

    def _partition_tensor(self, tensor):
        partition = partition_uniform(tensor.numel(), self.num_parts)
        start = partition[self.rank]
        length = partition[self.rank + 1] - start
        return tensor.detach().contiguous().view(-1).narrow(0, start, length).clone(), partition

    def full(self, device=None):
        device = device or self.orig_device
        numel = 1
        for s in self.orig_size:
            numel *= s
        flat = torch.zeros(numel, dtype=self.local_data.dtype, device=device)
        parts = []
        for r in range(self.num_parts):
            parts.append(flat.narrow(0, self.partition[r], self.partition[r + 1] - self.partition[r]))
        maxlen = max(p.numel() for p in parts)
        pad = torch.zeros(maxlen, dtype=flat.dtype, device=device)
        pad[: self.local_data.numel()] = self.local_data.to(device)
        gathered = [torch.zeros_like(pad) for _ in range(self.num_parts)]
        dist.all_gather(gathered, pad, group=self.group)
        for p, g in zip(parts, gathered):
            p.copy_(g[: p.numel()])
        return flat.view(self.orig_size)

    def to_meta(self):
        meta = [len(self.orig_size)] + list(self.orig_size) + [self.num_parts] + list(self.partition)
        return torch.LongTensor(meta).to(self.orig_device)

    def data(self):
        return self.local_data

    def local_size(self):
        return self.local_data.size()


def see_memory_usage(message, force=False):
    if not force:
        return
    if dist.is_initialized() and dist.get_rank() != 0:
        return
    gc.collect()
    acc = get_accelerator()
    gb = 1024 ** 3
    logger.info(message)
    logger.info(f"MA {acc.memory_allocated() / gb:.2f} GB  Max_MA {acc.max_memory_allocated() / gb:.2f} GB  "
                f"CA {acc.memory_reserved() / gb:.2f} GB  Max_CA {acc.max_memory_reserved() / gb:.2f} GB")
    vm = psutil.virtual_memory()
    logger.info(f"CPU Virtual Memory:  used = {(vm.total - vm.available) / gb:.2f} GB, percent = {vm.percent}%")
    acc.reset_peak_memory_stats()


def call_to_str(base, *args, **kwargs):
    parts = [repr(a) for a in args] + [f"{k}={v!r}" for k, v in kwargs.items()]
    return f"{base}({', '.join(parts)})"


def get_only_unique_item(items):
    s = set(items)
    if len(s) != 1:
        raise RuntimeError(f"expected there to be only one unique element in {items}")
    return next(iter(s))


def all_gather_dp_groups(groups_flat, partitioned_param_groups, dp_process_group, start_alignment_factor=1,
                         allgather_bucket_size=None):
    """In-place refresh of flat bit16 buffers from per-rank partitions."""
    for gid, flat in enumerate(groups_flat):
        pg = dp_process_group[gid] if isinstance(dp_process_group, (list, tuple)) else dp_process_group
        r = dist.get_rank(group=pg)
        dist.all_gather_into_tensor(flat, partitioned_param_groups[gid][r], group=pg)


def align_dense_tensors(tensor_list, alignment):
    n = sum(t.numel() for t in tensor_list)
    rem = n % alignment
    if rem:
        pad = torch.zeros(alignment - rem, device=tensor_list[0].device, dtype=tensor_list[0].dtype)
        return tensor_list + [pad]
    return tensor_list


def empty_cache():
    get_accelerator().empty_cache()
