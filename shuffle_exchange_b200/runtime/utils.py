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

    def full_size(self):
        return self.orig_size


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


# ================================================================ further reference utilities (``runtime/utils.py``)
class DummyOptim:
    """Stand-in optimizer for engines built without one (inference-style ZeRO-3): holds the parameters only."""

    def __init__(self, params):
        self.param_groups = [{"params": list(params)}]


graph_cache = {}


def graph_process(replay_first_step, func, *args, **kwargs):
    """Capture ``func(*args, **kwargs)`` into a CUDA graph on the first call (keyed by ``func.__name__``) and replay it
    afterwards.  ``replay_first_step``: also replay right after capture (capture itself does not execute)."""
    name = func.__name__
    if name not in graph_cache:
        acc = get_accelerator()
        stream = acc.Stream()
        stream.wait_stream(acc.current_stream())
        with acc.stream(stream):
            func(*args, **kwargs)              # warm-up outside capture
        acc.current_stream().wait_stream(stream)
        graph_cache[name] = acc.create_graph()
        with acc.capture_to_graph(graph_cache[name]):
            func(*args, **kwargs)
        if replay_first_step:
            acc.replay_graph(graph_cache[name])
    else:
        get_accelerator().replay_graph(graph_cache[name])


def noop_context(fn, *args, **kwargs):
    """Context manager that calls ``fn`` on entry and does nothing else."""
    import contextlib

    @contextlib.contextmanager
    def ctx():
        yield fn(*args, **kwargs) if callable(fn) else None
    return ctx()


def copy_to_device(item, device, criterion_func):
    """Deep-copy a nested structure, cloning onto ``device`` every element ``criterion_func`` selects."""
    if criterion_func(item):
        return item.to(device)
    if isinstance(item, list):
        return [copy_to_device(v, device, criterion_func) for v in item]
    if isinstance(item, tuple):
        return tuple(copy_to_device(v, device, criterion_func) for v in item)
    if isinstance(item, dict):
        return {k: copy_to_device(v, device, criterion_func) for k, v in item.items()}
    return item


def move_to_device(item, device, criterion_func=None):
    """Like :func:`copy_to_device` but in place for tensors (``.data`` re-pointed) — no second copy stays alive."""
    crit = criterion_func or torch.is_tensor
    if crit(item):
        moved = item.to(device)
        item.data = moved.data
        return item
    if isinstance(item, list):
        return [move_to_device(v, device, criterion_func) for v in item]
    if isinstance(item, tuple):
        return tuple(move_to_device(v, device, criterion_func) for v in item)
    if isinstance(item, dict):
        return {k: move_to_device(v, device, criterion_func) for k, v in item.items()}
    return item


def mask_nan_or_inf_with_val_inplace(input, device=None, val=-1.0):
    """Replace a non-finite scalar / tensor by ``val`` without a host sync (arithmetic select)."""
    bad = ~torch.isfinite(input)
    input.masked_fill_(bad, val) if input.dim() else input.copy_(torch.where(bad, torch.full_like(input, val), input))
    return input


def get_flattened_grad_norm(parameters, norm_type=2, mpu=None, grad_norm_mask=None):
    """Norm of the gradients of ``parameters`` (optionally masking index ranges per parameter), reduced over the
    model-parallel group; -1 when not finite."""
    parameters = [p for p in ([parameters] if torch.is_tensor(parameters) else parameters) if p.grad is not None]
    if not parameters:
        return 0.0
    grads = []
    for i, p in enumerate(parameters):
        g = p.grad.detach().float().reshape(-1)
        if grad_norm_mask is not None and len(grad_norm_mask[i]) > 0:
            keep = torch.ones_like(g, dtype=torch.bool)
            for lo, hi in grad_norm_mask[i].tolist():
                keep[lo:hi] = False
            g = g * keep
        grads.append(g)
    return get_global_norm_of_tensors(grads, norm_type=norm_type, mpu=mpu)


def get_grad_zeros(parameters, mpu=None):
    """Number of exactly-zero gradient elements over ``parameters`` (summed across model-parallel ranks)."""
    parameters = [p for p in ([parameters] if torch.is_tensor(parameters) else parameters) if p.grad is not None]
    total = 0.0
    tp_rank = bwc_rank(mpu)
    for p in parameters:
        if hasattr(p, "ds_pipe_replicated") and p.ds_pipe_replicated:
            continue
        if tp_rank > 0 and not is_model_parallel_parameter(p):
            continue
        total += float(p.grad.numel() - torch.count_nonzero(p.grad))
    t = torch.tensor([total], dtype=torch.float32, device=get_accelerator().current_device_name()
                     if get_accelerator().is_available() else "cpu")
    if mpu is not None and dist.is_initialized():
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=mpu.get_model_parallel_group())
    return t[0].item()


def bwc_rank(mpu) -> int:
    if mpu is None:
        return 0
    for n in ("get_tensor_model_parallel_rank", "get_model_parallel_rank"):
        if hasattr(mpu, n):
            return getattr(mpu, n)()
    return 0


def get_weight_norm(parameters, norm_type=2, mpu=None):
    """Norm of the parameter VALUES (replicated parameters counted once across model-parallel ranks)."""
    parameters = [parameters] if torch.is_tensor(parameters) else list(parameters)
    tp_rank = bwc_rank(mpu)
    own = [p.detach().float() for p in parameters
           if not (hasattr(p, "ds_pipe_replicated") and p.ds_pipe_replicated)
           and (tp_rank == 0 or is_model_parallel_parameter(p))]
    if not own:
        return 0.0
    return get_global_norm_of_tensors(own, norm_type=norm_type, mpu=mpu)


mem_alloced = 0
mem_cached = 0


def memory_status(msg, print_rank=-1, reset_max=False):
    """Allocated / cached device memory with deltas since the previous call."""
    global mem_alloced, mem_cached
    acc = get_accelerator()
    if not acc.is_available():
        return
    rank = dist.get_rank() if dist.is_initialized() else 0
    if print_rank != -1 and rank != print_rank:
        return
    acc.synchronize()
    if reset_max:
        acc.reset_max_memory_cached(); acc.reset_max_memory_allocated()
    new_alloced, new_cached = acc.memory_allocated(), acc.memory_cached()
    da, dc = new_alloced - mem_alloced, new_cached - mem_cached
    mem_alloced, mem_cached = new_alloced, new_cached
    gb = 1024 ** 3
    print(f"RANK={rank} MEMSTATS", msg, f"device={acc.current_device_name()} current alloc={new_alloced / gb:0.4f}GB "
          f"(delta={da / gb:0.4f}GB max={acc.max_memory_allocated() / gb:0.4f}GB) current cache={new_cached / gb:0.4f}GB "
          f"(delta={dc / gb:0.4f}GB max={acc.max_memory_cached() / gb:0.4f}GB)")


def get_ma_status():
    """Bytes currently allocated on the device (0 on other ranks than 0 when distributed, as the reference)."""
    if dist.is_initialized() and dist.get_rank() != 0:
        return 0
    return get_accelerator().memory_allocated() if get_accelerator().is_available() else 0


def all_gather_into_tensor_dp_groups(groups_flat, outputs, dp_process_group):
    """``outputs[i]`` := all-gather of ``groups_flat[i]`` over that group's data-parallel process group."""
    for gid, flat in enumerate(groups_flat):
        pg = dp_process_group[gid] if isinstance(dp_process_group, (list, tuple)) else dp_process_group
        dist.all_gather_into_tensor(outputs[gid], flat, group=pg)


def get_tensor_bytes(item) -> int:
    """Bytes of every tensor inside a nested list / tuple / dict structure."""
    if torch.is_tensor(item):
        return item.numel() * item.element_size()
    if isinstance(item, (list, tuple)):
        return sum(get_tensor_bytes(v) for v in item)
    if isinstance(item, dict):
        return sum(get_tensor_bytes(v) for v in item.values())
    return 0


def get_checkpoint_folder_size(save_dir, tag, local_rank=None) -> float:
    """Size in GB of ``save_dir/tag``; ``local_rank`` restricts the answer to local rank 0 (node-local storage)."""
    import os
    if local_rank is not None and local_rank != 0:
        return 0.0
    folder = os.path.join(save_dir, str(tag))
    total = 0
    for root, _, files in os.walk(folder):
        for f in files:
            fp = os.path.join(root, f)
            if not os.path.islink(fp):
                total += os.path.getsize(fp)
    return total / (1024 ** 3)


class TLinear(torch.nn.Linear):
    """Linear layer that stores its weight TRANSPOSED (``[in, out]``) so a ZeRO-3 gathered weight can be used by
    ``matmul`` without a transpose copy (reference ``runtime/utils.py:1110-1133``)."""

    def __init__(self, orig_layer, name=""):
        self.name = name
        super().__init__(orig_layer.weight.shape[1], orig_layer.weight.shape[0], bias=(orig_layer.bias is not None))
        self.weight.data = orig_layer.weight.data.t().contiguous()
        self.bias = orig_layer.bias
        self._fwd_func = self._fwd_bias_add if self.bias is not None else self._fwd

    def _fwd(self, input):
        if torch.is_grad_enabled() and (input.requires_grad or self.weight.requires_grad):
            return torch.matmul(input, self.weight)      # autograd path (the front end below records no graph)
        from ..ops import functional as _F
        return _F.matmul_nn(input, self.weight)          # inference: tcgen05 GEMM for bf16 CUDA operands

    def _fwd_bias_add(self, input):
        return self._fwd(input) + self.bias

    def forward(self, input):
        return self._fwd_func(input)


def get_inactive_params(param_list):
    """ZeRO-3 parameters that are currently partitioned away (must be gathered before use)."""
    from .zero.partition_parameters import ZeroParamStatus
    return [p for p in param_list if hasattr(p, "ds_id") and p.ds_status == ZeroParamStatus.NOT_AVAILABLE]


def get_norm_with_moe_layers(non_expert_norm, mpu, expert_tensors, norm_type=2):
    """Global gradient norm of a model with experts: non-expert norm (already global) combined with every expert
    group's norm, each reduced over ITS expert-parallel group (experts are not replicated across it)."""
    def to_tensor(v):
        return v.detach().float().reshape(1) if torch.is_tensor(v) else torch.tensor([float(v)])
    dev = next((t.device for ts in expert_tensors.values() for t in ts), torch.device("cpu"))
    inf_norm = norm_type == float("inf")
    total = to_tensor(non_expert_norm).to(dev)
    total = total if inf_norm else total ** norm_type
    for group_name, tensors in expert_tensors.items():
        if not tensors:
            continue
        if inf_norm:
            local = torch.stack([t.detach().abs().max().float() for t in tensors]).max().reshape(1)
        else:
            local = torch.stack([t.detach().float().norm(norm_type) ** norm_type for t in tensors]).sum().reshape(1)
        group = groups._get_expert_parallel_group(group_name) if dist.is_initialized() else None
        if group is not None:
            dist.all_reduce(local, op=dist.ReduceOp.MAX if inf_norm else dist.ReduceOp.SUM, group=group)
        total = torch.maximum(total, local) if inf_norm else total + local
    total = total if inf_norm else total ** (1.0 / norm_type)
    out = float(total.item())
    return -1 if (out == float("inf") or out != out) else out


def get_norm_with_moe_layers_fast(all_groups_norm, group):
    """Variant for BF16 / ZeRO-1 where ``all_groups_norm`` already holds this rank's contribution: average the scaled
    norm over ``group`` so expert parameters (sharded over expert-data-parallel ranks) are counted once."""
    scaled = all_groups_norm * 1.0 / float(dist.get_world_size(group=group))
    t = torch.tensor([float(scaled)], dtype=torch.float32,
                     device=get_accelerator().current_device_name() if get_accelerator().is_available() else "cpu")
    dist.all_reduce(t, group=group)
    return t.item()


def offload_adam_states(optimizer, device, pin_memory=False, non_blocking=False):
    from .zero.offload_states import offload_adam_states as _f
    return _f(optimizer, device, pin_memory=pin_memory, non_blocking=non_blocking)


def reload_adam_states(optimizer, device, non_blocking=False):
    from .zero.offload_states import reload_adam_states as _f
    return _f(optimizer, device, non_blocking=non_blocking)


def compare_tensors_in_structures(inputs1, inputs2) -> bool:
    """Same nesting, same keys, and ``torch.equal`` tensors (non-tensor leaves compared with ``==``)."""
    if type(inputs1) != type(inputs2):
        return False
    if isinstance(inputs1, (list, tuple)):
        return len(inputs1) == len(inputs2) and all(compare_tensors_in_structures(a, b) for a, b in zip(inputs1, inputs2))
    if isinstance(inputs1, dict):
        return inputs1.keys() == inputs2.keys() and all(compare_tensors_in_structures(inputs1[k], inputs2[k])
                                                        for k in inputs1)
    if torch.is_tensor(inputs1):
        return inputs1.device == inputs2.device and torch.equal(inputs1, inputs2)
    return inputs1 == inputs2
