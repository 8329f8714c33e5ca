"""Token / position embedding tables (reference: ``common_parameters/embedding_parameters.py``)."""
import torch

from ..parameter_base import ParameterBase


class EmbeddingParameter(ParameterBase):
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.params
