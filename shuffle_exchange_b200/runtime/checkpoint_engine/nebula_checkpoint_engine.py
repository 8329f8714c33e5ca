"""Nebula (Azure tiered checkpoint service) engine — reference: ``runtime/checkpoint_engine/nebula_checkpoint_engine.py``.
The service SDK (``torch_nebula``) is proprietary and not present in this image; constructing the engine without it
raises ``ImportError`` which ``create_checkpoint_engine`` turns into a fallback to ``torch.save`` exactly as the
reference does (utils.py:30-36)."""
import os

import torch

from ...utils.logging import log_dist, logger
from .checkpoint_engine import CheckpointEngine


class NebulaCheckpointEngine(CheckpointEngine):
    def __init__(self, config_params=None):
        super().__init__(config_params)
        import torch_nebula  # noqa: F401  (raises ImportError when the SDK is absent)
        self.nebula = torch_nebula
        self.checkpoint = None
        self.tag_flag = None
        self.enable_nebula_load = config_params.get("enable_nebula_load", True)
        self.nebula_load_path = config_params.get("load_path") or config_params.get("persistent_storage_path")
        self.nebula.init(persistent_storage_path=config_params.get("persistent_storage_path"),
                         persistent_time_interval=config_params.get("persistent_time_interval", 100),
                         num_of_version_in_retention=config_params.get("num_of_version_in_retention", 2))

    def create(self, info):
        log_dist(f"[Nebula] Start Checkpoint for tag:{info}", ranks=[0])
        self.checkpoint = self.nebula.Checkpoint(info, -2)

    def save(self, state_dict, path: str):
        tag = os.path.basename(os.path.dirname(path))
        self.checkpoint.save(os.path.basename(path), state_dict)

    def load(self, path: str, map_location=None):
        if not self.enable_nebula_load:
            return torch.load(path, map_location=map_location, weights_only=False)
        tag = os.path.basename(os.path.dirname(path))
        ckpt = self.nebula.get_checkpoint(tag=tag, persist_path=self.nebula_load_path) or self.nebula.get_latest_checkpoint(persist_path=self.nebula_load_path)
        return ckpt.load(os.path.basename(path), map_location=map_location)

    def commit(self, info):
        self.checkpoint.commit()
        log_dist(f"[Nebula] all files for {info} are saved in tier1; persisting asynchronously", ranks=[0])
        return True
