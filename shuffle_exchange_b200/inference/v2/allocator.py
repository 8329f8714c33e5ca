"""Tiny allocation helpers (reference ``inference/v2/allocator.py``): ``on_device`` marks a tensor as resident on the
current device, ``empty_from`` carves typed views out of a pre-allocated byte/flat buffer so steady-state inference
never calls the allocator."""
from __future__ import annotations

from functools import reduce
from typing import Iterable

import torch


def _dev():
    return torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")


class Allocator:
    cache = {}

    @staticmethod
    def empty_from(tensor: torch.Tensor, shape: Iterable[int]) -> torch.Tensor:
        shape = tuple(int(s) for s in shape)
        n = reduce(lambda a, b: a * b, shape, 1)
        if n == 0:
            raise ValueError("Cannot create empty tensor with size 0")
        if n > tensor.numel():
            raise ValueError(f"view of {n} elements does not fit the backing buffer of {tensor.numel()}")
        key = (tensor.data_ptr(), tensor.dtype, shape)
        view = Allocator.cache.get(key)
        if view is None or view.data_ptr() != tensor.data_ptr():
            view = tensor.flatten()[:n].view(shape)
            Allocator.cache[key] = view
        return view


empty_from = Allocator.empty_from


def on_device(method):
    """Decorator: tensors returned by ``method`` are moved to the current device (non-tensors pass through)."""

    def wrapped(self, *args, **kwargs):
        out = method(self, *args, **kwargs)
        if isinstance(out, torch.Tensor):
            return out.to(_dev())
        return out

    return wrapped
