"""Guards for code that must not be traced by ``torch.compile`` (reference ``deepspeed/runtime/compiler.py:9-71``).
The framework itself never relies on a tracing compiler; these helpers exist so user code decorated with them keeps
working and so hooks that do Python-side bookkeeping are skipped cleanly if a user compiles their module."""
from __future__ import annotations

import functools

import torch

from ..utils.torch import required_torch_version

try:
    from torch.compiler import is_compiling as _torch_is_compiling
except ImportError:      # pragma: no cover - very old torch
    try:
        from torch._dynamo.external_utils import is_compiling as _torch_is_compiling
    except ImportError:
        _torch_is_compiling = lambda: False  # noqa: E731


def is_compile_supported() -> bool:
    return required_torch_version(min_version=2.1)


def disable(func):
    """Exclude ``func`` from tracing."""
    return torch.compiler.disable(func) if is_compile_supported() else func


def enable(min_version=None):
    """``@enable`` / ``@enable("2.7.0")``: allow tracing of the function only from that torch version on."""

    def decorator(func):
        guarded = None

        @functools.wraps(func)
        def wrapper(*args, **kwargs):
            nonlocal guarded
            if min_version is None or required_torch_version(min_version=min_version):
                return func(*args, **kwargs)
            if guarded is None:
                guarded = disable(func)
            return guarded(*args, **kwargs)

        return wrapper

    if callable(min_version):      # bare @enable
        func, min_version = min_version, None
        return decorator(func)
    return decorator


def is_compiling() -> bool:
    return _torch_is_compiling()
