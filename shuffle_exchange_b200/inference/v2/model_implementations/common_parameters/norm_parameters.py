"""Normalisation scale / shift (reference: ``common_parameters/norm_parameters.py``)."""
import torch

from ..parameter_base import ParameterBase


class NormParameter(ParameterBase):
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.params
