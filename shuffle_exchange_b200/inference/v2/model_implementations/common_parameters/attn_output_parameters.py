"""Attention output projection (reference: ``common_parameters/attn_output_parameters.py``)."""
import torch

from ..parameter_base import ParameterBase


class AttentionOutputParameter(ParameterBase):
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.params
