"""Rotary inverse frequencies stored in a checkpoint (reference: ``common_parameters/invfreq_parameters.py``); the
attention kernels derive them from ``rope_theta``, so the tensor is only kept for models that ship a non-standard
table."""
import torch

from ..parameter_base import ParameterBase


class InvFreqParameter(ParameterBase):
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.params.float()
