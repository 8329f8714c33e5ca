"""Move optimizer/model state between device and (pinned) host memory on demand.

Reference: ``runtime/zero/offload_states.py:14-87`` + ``stage3.py:2960-3106`` (``offload_states`` /
``reload_states`` with ``OffloadStateTypeEnum`` include-lists, ``pin_memory`` and ``non_blocking``).  Works on any of
this framework's sharded optimizers: the states are the per-flat-group tensors (fp32 master shard, optimizer
moments, low-precision shard, gradient shard), so one ``.to()`` per flat group moves a whole layer's state.
"""
from __future__ import annotations

from enum import Enum
from typing import Iterable, Optional, Set

import torch


class OffloadDeviceEnum(str, Enum):
    none = "none"
    cpu = "cpu"
    nvme = "nvme"


class OffloadStateTypeEnum(str, Enum):
    optim_states = "optim_states"
    hp_params = "hp_params"
    lp_params = "lp_params"
    lp_grads = "lp_grads"
    contiguous_grad_buffer = "contiguous_grad_buffer"


def _to(t: torch.Tensor, device, pin_memory: bool, non_blocking: bool) -> torch.Tensor:
    if t is None or t.device.type == torch.device(device).type:
        return t
    if torch.device(device).type == "cpu":
        host = torch.empty(t.shape, dtype=t.dtype, device="cpu", pin_memory=pin_memory and torch.cuda.is_available())
        host.copy_(t, non_blocking=non_blocking)
        return host
    return t.to(device, non_blocking=non_blocking)


def offload_adam_states(optimizer, device, pin_memory: bool = False, non_blocking: bool = False):
    """Move ``exp_avg`` / ``exp_avg_sq`` (and any other tensor state) of a torch-style optimizer."""
    for state in optimizer.state.values():
        for k, v in list(state.items()):
            if torch.is_tensor(v) and v.numel() > 1:
                state[k] = _to(v, device, pin_memory, non_blocking)


def reload_adam_states(optimizer, device, non_blocking: bool = False):
    offload_adam_states(optimizer, device, pin_memory=False, non_blocking=non_blocking)


def _wanted(include: Optional[Iterable[OffloadStateTypeEnum]], kind: OffloadStateTypeEnum) -> bool:
    return include is None or kind in set(include)


def offload_optimizer_states(zero_optimizer, include=None, device=OffloadDeviceEnum.cpu, pin_memory=True,
                             non_blocking=False):
    """Offload the selected state kinds of a sharded optimizer; records what moved so ``reload`` can undo it."""
    dev = "cpu" if str(getattr(device, "value", device)) == "cpu" else device
    moved: Set[OffloadStateTypeEnum] = getattr(zero_optimizer, "_offloaded_states", set())
    groups = getattr(zero_optimizer, "groups", [])
    if _wanted(include, OffloadStateTypeEnum.optim_states):
        offload_adam_states(zero_optimizer.optimizer, dev, pin_memory, non_blocking)
        moved.add(OffloadStateTypeEnum.optim_states)
    if _wanted(include, OffloadStateTypeEnum.hp_params):
        for fg in groups:
            fg.hp_param.data = _to(fg.hp_param.data, dev, pin_memory, non_blocking)
        moved.add(OffloadStateTypeEnum.hp_params)
    if _wanted(include, OffloadStateTypeEnum.lp_params):
        for fg in groups:
            if getattr(fg, "_symm_shard", None) is None:          # symmetric-heap shards must stay registered
                fg.lp_shard = _to(fg.lp_shard, dev, pin_memory, non_blocking)
        moved.add(OffloadStateTypeEnum.lp_params)
    if _wanted(include, OffloadStateTypeEnum.lp_grads):
        for fg in groups:
            if getattr(fg, "grad_shard", None) is not None:
                fg.grad_shard = _to(fg.grad_shard, dev, pin_memory, non_blocking)
        moved.add(OffloadStateTypeEnum.lp_grads)
    if _wanted(include, OffloadStateTypeEnum.contiguous_grad_buffer):
        for fg in groups:
            if getattr(fg, "flat_grad", None) is not None and getattr(fg, "_symm_grad", None) is None:
                fg.release_flat_grad()
        moved.add(OffloadStateTypeEnum.contiguous_grad_buffer)
    zero_optimizer._offloaded_states = moved
    if not non_blocking and torch.cuda.is_available():
        torch.cuda.synchronize()
    if torch.cuda.is_available():
        torch.cuda.empty_cache()


def reload_optimizer_states(zero_optimizer, non_blocking=False):
    moved = getattr(zero_optimizer, "_offloaded_states", set())
    if not moved:
        return
    dev = zero_optimizer.groups[0].device if getattr(zero_optimizer, "groups", None) else "cuda"
    if OffloadStateTypeEnum.optim_states in moved:
        reload_adam_states(zero_optimizer.optimizer, dev, non_blocking)
    for fg in getattr(zero_optimizer, "groups", []):
        if OffloadStateTypeEnum.hp_params in moved:
            fg.hp_param.data = fg.hp_param.data.to(dev, non_blocking=non_blocking)
        if OffloadStateTypeEnum.lp_params in moved and fg.lp_shard.device.type == "cpu":
            fg.lp_shard = fg.lp_shard.to(dev, non_blocking=non_blocking)
        if OffloadStateTypeEnum.lp_grads in moved and getattr(fg, "grad_shard", None) is not None:
            fg.grad_shard = fg.grad_shard.to(dev, non_blocking=non_blocking)
    zero_optimizer._offloaded_states = set()
    if not non_blocking and torch.cuda.is_available():
        torch.cuda.synchronize()


def get_state_devices(model, state: OffloadStateTypeEnum) -> Set[torch.device]:
    """Devices currently holding one kind of state of ``model`` (a training engine): what tests and users check after
    ``engine.offload_states`` / ``reload_states`` (reference ``offload_states.py:49-72``).  States live per flat group
    here, so the answer is one device per group rather than one per parameter."""
    opt = model.optimizer
    groups = getattr(opt, "groups", [])
    state = OffloadStateTypeEnum(getattr(state, "value", state))
    if state == OffloadStateTypeEnum.hp_params:
        return {fg.hp_param.device for fg in groups}
    if state == OffloadStateTypeEnum.lp_params:
        return {fg.lp_shard.device for fg in groups}
    if state == OffloadStateTypeEnum.lp_grads:
        return {fg.grad_shard.device for fg in groups if getattr(fg, "grad_shard", None) is not None}
    if state == OffloadStateTypeEnum.contiguous_grad_buffer:
        return {fg.flat_grad.device for fg in groups if getattr(fg, "flat_grad", None) is not None}
    inner = getattr(opt, "optimizer", opt)
    return {v.device for st in inner.state.values() for v in st.values() if torch.is_tensor(v) and v.numel() > 1}
