"""``torch.autocast`` integration: which parameters may be communicated / gathered in the autocast dtype.

Reference: ``deepspeed/runtime/torch_autocast.py:12-100``.  With autocast the model keeps fp32 parameters; modules on
the "lower-precision safe" list (Linear / Conv by default) get ``param.autocast_dtype`` so ZeRO gathers and reduces
those parameters in bf16/fp16 while everything else stays fp32."""
from __future__ import annotations

import importlib
from typing import Iterable, List, Optional, Set

import torch

from ..utils.logging import logger

LOWER_PRECISION_SAFE_MODULES = [torch.nn.Linear, torch.nn.Conv1d, torch.nn.Conv2d, torch.nn.Conv3d]
TORCH_AUTOCAST_INITIALIZED = False
_WARNED_NESTED_AUTOCAST = False


def _validate_auto_cast_settings(engine) -> None:
    assert not engine.fp16_enabled(), "Cannot enable both torch autocast and fp16"
    assert not engine.bfloat16_enabled(), "Cannot enable both torch autocast and bfloat16"
    if hasattr(engine, "zero_quantized_weights"):
        assert not engine.zero_quantized_weights(), "Cannot enable both torch autocast and zero quantized weights"
    assert all(p.dtype == torch.float32 for p in engine.module.parameters()), \
        "All parameters must be float32 for torch autocast"
    assert engine.communication_data_type == torch.float32, \
        "Communication data type must be float32 for torch autocast"


def _resolve(names: Optional[List[str]]):
    if names is None:
        return list(LOWER_PRECISION_SAFE_MODULES)
    classes = []
    for dotted in names:
        try:
            pkg, cls = dotted.rsplit(".", 1)
            classes.append(getattr(importlib.import_module(pkg), cls))
        except Exception as e:
            raise ValueError(f"Failed to import lower precision safe module {dotted}: {e}")
    return classes


def init_autocast_params(engine, dtype: torch.dtype,
                         torch_autocast_lower_precision_safe_modules: Optional[List[str]]) -> None:
    global TORCH_AUTOCAST_INITIALIZED
    _validate_auto_cast_settings(engine)
    safe = tuple(_resolve(torch_autocast_lower_precision_safe_modules))
    for module in engine.module.modules():
        if type(module) in safe:
            for p in module.parameters(recurse=False):
                p.autocast_dtype = dtype
    TORCH_AUTOCAST_INITIALIZED = True


def is_autocast_initialized() -> bool:
    return TORCH_AUTOCAST_INITIALIZED


def get_default_autocast_lower_precision_modules() -> List[str]:
    return [f"{c.__module__}.{c.__name__}" for c in LOWER_PRECISION_SAFE_MODULES]


def get_autocast_dtype(param: torch.nn.Parameter) -> torch.dtype:
    return getattr(param, "autocast_dtype", param.dtype)


def has_autocast_dtype(param: torch.nn.Parameter) -> bool:
    return hasattr(param, "autocast_dtype")


def get_all_autocast_dtypes(params: Iterable) -> Set[torch.dtype]:
    return {get_autocast_dtype(p) for p in params}


def sort_dtypes(dtypes: List[torch.dtype]) -> List[torch.dtype]:
    return sorted(dtypes, key=str)


def validate_nested_autocast(engine) -> None:
    """A user-side ``torch.autocast`` is only legal when the engine was configured for autocast too (otherwise the
    communication dtype chosen at init would be wrong); when both are on, warn once that the outer one is redundant."""
    global _WARNED_NESTED_AUTOCAST
    active = torch.is_autocast_enabled() or (hasattr(torch, "is_autocast_cpu_enabled")
                                              and torch.is_autocast_cpu_enabled())
    if not active:
        return
    if engine.torch_autocast_enabled():
        if not _WARNED_NESTED_AUTOCAST:
            logger.warning("torch.autocast context detected outside the engine; it is unnecessary when autocast is "
                           "enabled through the config.")
            _WARNED_NESTED_AUTOCAST = True
        return
    raise AssertionError("torch.autocast is enabled outside the engine, but not in the config. Enable torch.autocast "
                         "through the config so the correct communication dtype is used.")
