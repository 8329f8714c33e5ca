"""Enumerations and small helpers of the v2 inference stack (reference ``inference/v2/inference_utils.py``)."""
from __future__ import annotations

from enum import Enum, IntEnum
from typing import Dict

import torch


class NormTypeEnum(Enum):
    LayerNorm = "layer_norm"
    RMSNorm = "rms_norm"


class DtypeEnum(Enum):
    """Values are tuples so several spellings parse to one member: ``DtypeEnum("bf16") is DtypeEnum(torch.bfloat16)``."""
    fp16 = (torch.float16, "torch.float16", "fp16", "float16", "half")
    fp32 = (torch.float32, "torch.float32", "fp32", "float32", "float")
    bf16 = (torch.bfloat16, "torch.bfloat16", "bf16", "bfloat16", "bfloat")
    int8 = (torch.int8, "torch.int8", "int8")

    @classmethod
    def _missing_(cls, value):
        for m in cls:
            if value in m.value:
                return m
        return None

    @staticmethod
    def from_str(value: str) -> "DtypeEnum":
        m = DtypeEnum(value)
        if m is None:
            raise ValueError(f"'{value}' is not a valid DtypeEnum")
        return m

    def __repr__(self):
        return str(self.value[0])

    @property
    def torch(self) -> torch.dtype:
        return self.value[0]


ELEM_SIZES: Dict[torch.dtype, int] = {torch.float16: 2, torch.bfloat16: 2, torch.float32: 4, torch.float64: 8,
                                      torch.int8: 1, torch.uint8: 1, torch.int16: 2, torch.int32: 4,
                                      torch.int64: 8, torch.bool: 1}


def elem_size(dtype: torch.dtype) -> int:
    try:
        return ELEM_SIZES[dtype]
    except KeyError:
        raise ValueError(f"Unknown dtype size for {dtype}")


class ActivationType(IntEnum):
    """Order matters: the gated variants sit right after their plain counterparts (``is_gated`` relies on names)."""
    GELU = 0
    RELU = 1
    SILU = 2
    GEGLU = 3
    ReGLU = 4
    SiGLU = 5
    IDENTITY = 6
    InvalidType = -1


def is_gated(act_fn: ActivationType) -> bool:
    if not isinstance(act_fn, ActivationType):
        act_fn = ActivationType(act_fn)
    return act_fn in (ActivationType.GEGLU, ActivationType.ReGLU, ActivationType.SiGLU)


def ceil_div(a: int, b: int) -> int:
    return -(-a // b)
