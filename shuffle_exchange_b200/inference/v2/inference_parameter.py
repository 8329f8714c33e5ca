"""``InferenceParameter`` — a tensor that carries named companion tensors (quantisation scales, zero points, ...).

The companions live in one dict on the instance and are reachable as attributes; moving the parameter with ``.to()``
moves them along, re-wrapping a plain tensor keeps them.  Same constructor surface as the reference class
(``deepspeed/inference/v2/inference_parameter.py``): ``initialize(core, **aux)``, ``initialize_raw(**{CORE_PARAM: core,
...})``, ``aux_attrs``."""
from __future__ import annotations

from typing import Dict

import torch

CORE_PARAM = "_ds_core_param_key"
STR_TO_DTYPE = {str(dt): dt for dt in (torch.float32, torch.float64, torch.float16, torch.bfloat16, torch.int64,
                                       torch.int32, torch.int16, torch.int8, torch.uint8, torch.bool)}
_SLOT = "_aux_attrs"


def _adopt(dst: torch.Tensor, companions: Dict[str, torch.Tensor]) -> None:
    object.__setattr__(dst, _SLOT, companions)
    for name, t in companions.items():
        object.__setattr__(dst, name, t)


class InferenceParameter(torch.Tensor):

    @staticmethod
    def __new__(cls, tensor, *args, **kwargs):
        self = super().__new__(cls, tensor, *args, **kwargs)
        inherited = getattr(tensor, _SLOT, None)
        if inherited is not None:
            _adopt(self, inherited)
        return self

    @property
    def aux_attrs(self) -> Dict[str, torch.Tensor]:
        return getattr(self, _SLOT, {})

    def to(self, *args, **kwargs):
        moved = super().to(*args, **kwargs)
        companions = dict(self.aux_attrs)
        # companions follow DEVICE moves only: a dtype cast of the payload must not touch fp32 scales
        target = kwargs.get("device", args[0] if args and not isinstance(args[0], torch.dtype) else None)
        if target is not None and not isinstance(target, torch.Tensor):
            try:
                dev = torch.device(target)
                companions = {k: v.to(dev) for k, v in companions.items()}
            except (TypeError, RuntimeError):
                pass
        _adopt(moved, companions)
        return moved

    @classmethod
    def initialize(cls, core_param: torch.Tensor, **kwargs) -> "InferenceParameter":
        param = InferenceParameter(core_param)
        for name, t in kwargs.items():
            if not isinstance(t, torch.Tensor):
                raise ValueError(f"Attribute {name} must be a tensor.")
            if hasattr(param, name):
                raise ValueError(f"Attribute {name} already exists on param.")
        _adopt(param, dict(kwargs))
        return param

    @classmethod
    def initialize_raw(cls, **kwargs) -> "InferenceParameter":
        try:
            core = kwargs.pop(CORE_PARAM)
        except KeyError:
            raise ValueError(f"Must provide core parameter, with key {CORE_PARAM}.") from None
        return cls.initialize(core, **kwargs)
