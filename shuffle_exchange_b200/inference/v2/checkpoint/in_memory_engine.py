"""Serve parameters of an already instantiated ``nn.Module`` (reference ``checkpoint/in_memory_engine.py``)."""
from typing import Iterable, Tuple

import torch

from .base_engine import CheckpointEngineBase


class InMemoryModelEngine(CheckpointEngineBase):

    def __init__(self, model: torch.nn.Module) -> None:
        super().__init__()
        self.model = model
        self.model_config = getattr(model, "config", None)

    def parameters(self) -> Iterable[Tuple[str, torch.Tensor]]:
        seen = set()
        for name, t in self.model.state_dict().items():
            seen.add(name)
            yield name, t
        # tied output embedding: HF drops lm_head.weight from state_dict views in some versions
        if "lm_head.weight" not in seen and hasattr(self.model, "get_output_embeddings"):
            out = self.model.get_output_embeddings()
            if out is not None and getattr(out, "weight", None) is not None:
                yield "lm_head.weight", out.weight
