"""Output projection onto the vocabulary (reference: ``common_parameters/unembed_parameters.py``)."""
import torch

from ..parameter_base import ParameterBase


class UnembedParameter(ParameterBase):
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.params
