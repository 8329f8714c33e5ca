"""ZeRO helper functions (reference: ``runtime/zero/utils.py:20-170``)."""
from __future__ import annotations

import os
from typing import List

import torch

from ... import comm as dist
from ...utils.logging import logger


class ZeRORuntimeException(Exception):
    pass


def _initialize_parameter_parallel_groups(parameter_parallel_size=None):
    world = dist.get_world_size()
    size = parameter_parallel_size or world
    logger.info(f"data_parallel_size: {world}, parameter_parallel_size: {size}")
    assert world % size == 0, "world size should be divisible by parameter parallel size"
    rank, my_group = dist.get_rank(), None
    for i in range(world // size):
        ranks = list(range(i * size, (i + 1) * size))
        g = dist.new_group(ranks)
        if rank in ranks:
            my_group = g
    return my_group


ZERO_SUPPORTED_OPTIMIZERS: List[type] = [torch.optim.Adam, torch.optim.AdamW, torch.optim.SGD, torch.optim.Adagrad]


def _register_native_optimizers():
    from ...ops.adam import DeepSpeedCPUAdam, FusedAdam
    from ...ops.adagrad import DeepSpeedCPUAdagrad
    from ...ops.lion import DeepSpeedCPULion, FusedLion
    from ...ops.lamb import FusedLamb
    for cls in (FusedAdam, DeepSpeedCPUAdam, DeepSpeedCPUAdagrad, DeepSpeedCPULion, FusedLion, FusedLamb):
        if cls not in ZERO_SUPPORTED_OPTIMIZERS:
            ZERO_SUPPORTED_OPTIMIZERS.append(cls)


def is_zero_supported_optimizer(optimizer) -> bool:
    _register_native_optimizers()
    if dist.is_initialized() and dist.get_rank() == 0:
        logger.info(f"Checking ZeRO support for optimizer={optimizer.__class__.__name__} type={type(optimizer)}")
    return type(optimizer) in ZERO_SUPPORTED_OPTIMIZERS


def get_lst_from_rank0(lst: List[int]) -> List[int]:
    """Every rank returns rank 0's copy of ``lst`` (lengths must match)."""
    t = torch.tensor(lst if dist.get_rank() == 0 else [-1] * len(lst), dtype=torch.int64,
                     device=torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else "cpu")
    dist.broadcast(t, src=0, async_op=False)
    return t.cpu().tolist()


def assert_ints_same_as_other_ranks(ints: List[int]) -> None:
    """Guards against ranks taking different code paths (e.g. different fetch orders), which deadlocks collectives."""
    rank0 = get_lst_from_rank0(ints)
    if ints != rank0:
        raise RuntimeError(f"disagreement between rank0 and rank{dist.get_rank()}: rank0: {rank0}, rank{dist.get_rank()}: {ints}")


def is_builtin_type(obj) -> bool:
    return obj.__class__.__module__ in ("__builtin__", "builtins")


def isinstance_namedtuple(obj) -> bool:
    return isinstance(obj, tuple) and hasattr(obj, "_asdict") and hasattr(obj, "_fields")


def is_zero_param(parameter) -> bool:
    return torch.is_tensor(parameter) and (hasattr(parameter, "ds_id") or hasattr(parameter, "_sxb_link"))


def apply_to_tensors_only(function, value, warning_msg_fn=None):
    """Apply ``function`` to every tensor inside nested lists / tuples / dicts, preserving structure."""
    if isinstance(value, (tuple, list)):
        touched = [apply_to_tensors_only(function, v, warning_msg_fn) for v in value]
        if isinstance_namedtuple(value):
            return value.__class__(*touched)
        return value.__class__(touched)
    if isinstance(value, dict):
        return {k: apply_to_tensors_only(function, v, warning_msg_fn) for k, v in value.items()}
    if isinstance(value, torch.Tensor):
        return function(value)
    if not is_builtin_type(value) and warning_msg_fn is not None:
        global _warned
        if not _warned and (not dist.is_initialized() or dist.get_rank() == 0):
            logger.warning(warning_msg_fn(value))
            _warned = True
    return value


_warned = False


def get_mapping_to_flat_buffer(tensors: List[torch.Tensor]):
    """(tensor, offset, numel) triples for laying ``tensors`` out back to back."""
    out, off = [], 0
    for t in tensors:
        out.append((t, off, t.numel()))
        off += t.numel()
    return out
