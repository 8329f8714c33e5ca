"""NHWC bias add (+ optional second activation) used by the diffusers UNet path (reference
``ops/transformer/inference/bias_add.py``)."""
from typing import Optional

import torch


def nhwc_bias_add(activation: torch.Tensor, bias: torch.Tensor, other: Optional[torch.Tensor] = None,
                  other_bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``activation`` is channels-last ``[N, C, H, W]``; ``bias`` is ``[C]``.  ``out = activation + bias (+ other
    (+ other_bias))``."""
    out = activation + bias.view(1, -1, 1, 1)
    if other is not None:
        out = out + (other if other_bias is None else other + other_bias.view(1, -1, 1, 1))
    return out
