"""Mixture-of-experts parameters (reference: ``common_parameters/moe_parameters.py``)."""
import torch

from ..parameter_base import ParameterBase, ParamList


class MoEGatingWeightParameter(ParameterBase):
    """Router ``[n_experts, model_dim]``."""
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.params


class UnfusedMoEMLP1Parameter(ParameterBase):
    """Per-expert first projections -> ``[E, F, H]``."""
    experts: ParamList("n_experts")

    def finalize(self) -> torch.Tensor:
        return torch.stack(list(self.experts)).contiguous()


class UnfusedMoEMLP2Parameter(ParameterBase):
    """Per-expert second projections -> ``[E, H, F]``."""
    experts: ParamList("n_experts")

    def finalize(self) -> torch.Tensor:
        return torch.stack(list(self.experts)).contiguous()


class UnfusedMoEGatedMLPParameter(ParameterBase):
    """Per-expert gate and up projections -> ``[E, 2F, H]`` with gate|up halves."""
    gating_experts: ParamList("n_experts")
    up_experts: ParamList("n_experts")

    def finalize(self) -> torch.Tensor:
        return torch.stack([torch.cat([g, u], dim=0) for g, u in zip(self.gating_experts, self.up_experts)]).contiguous()
