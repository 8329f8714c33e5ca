"""Small helpers mirroring ``deepspeed/compile/util.py``."""
from __future__ import annotations

import torch

from . import backend as _backend


def is_deepcompile_supported() -> bool:
    """The schedule compiler needs nothing beyond a CUDA device (the reference needs torch>=2.6 + its native op)."""
    return True


class _Handle:
    """Stand-in for the reference's native ``dc`` handle: lifecycle calls the engine makes."""

    def cleanup(self):
        _backend._state.clear()
        _backend._schedule.clear()

    def start_backward(self, update: bool):
        _backend._state["in_backward"] = True
        _backend._state["is_boundary"] = bool(update)

    def end_backward(self):
        _backend._state["in_backward"] = False


_handle = _Handle()


def get_deepcompile_handle():
    return _handle


def deepcompile_backward_prologue(is_gradient_accumulation_boundary: bool):
    _handle.start_backward(is_gradient_accumulation_boundary)


def get_input_nodes(plan):
    return [n for n in plan.nodes if n.kind == "gather"]
