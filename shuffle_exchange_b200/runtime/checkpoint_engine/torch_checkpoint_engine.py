"""torch.save / torch.load engine (reference: torch_checkpoint_engine.py)."""
import torch

from ...utils.logging import log_dist
from .checkpoint_engine import CheckpointEngine


class TorchCheckpointEngine(CheckpointEngine):
    def create(self, info):
        log_dist(f"[Torch] Checkpoint {info} is about to be saved!", ranks=[0])

    def save(self, state_dict, path: str):
        torch.save(state_dict, path)

    def load(self, path: str, map_location=None):
        return torch.load(path, map_location=map_location, weights_only=False)

    def commit(self, info):
        log_dist(f"[Torch] Checkpoint {info} is ready now!", ranks=[0])
        return True
