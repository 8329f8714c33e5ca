"""A checkpoint engine yields ``(parameter_name, tensor)`` pairs, names in the HuggingFace convention of the
architecture (reference ``inference/v2/checkpoint/base_engine.py``)."""
from abc import ABC, abstractmethod
from typing import Iterable, Tuple

import torch

MEGATRON = "megatron"
HUGGINGFACE = "huggingface"


class CheckpointEngineBase(ABC):

    @abstractmethod
    def parameters(self) -> Iterable[Tuple[str, torch.Tensor]]:
        ...
