"""Dense MLP projections (reference: ``common_parameters/mlp_parameters.py``)."""
import torch

from ..parameter_base import ParameterBase


class MLP1Parameter(ParameterBase):
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.params


class GatedMLPParameter(ParameterBase):
    """gate and up projections stacked gate|up (the activation is applied to the first half)."""
    gate_params: torch.Tensor
    up_params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return torch.cat([self.gate_params, self.up_params], dim=0).contiguous()


class MLP2Parameter(ParameterBase):
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.params
