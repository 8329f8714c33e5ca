"""Checkpoint engine interface (reference: deepspeed/runtime/checkpoint_engine/checkpoint_engine.py:14-60)."""
import os
from dataclasses import dataclass


@dataclass
class CheckpointCommitInfo:
    """What has to happen once every file of a checkpoint is durable: which tag it was, where, and whether the
    ``latest`` marker moves to it."""
    tag: str
    save_dir: str
    save_latest: bool

    def __str__(self):
        return str(self.tag)


class CheckpointEngine:
    def __init__(self, config_params=None):
        self.config_params = config_params
        self.name = None

    def create(self, info):
        """Announce the start of a checkpoint identified by ``info`` (a tag or a :class:`CheckpointCommitInfo`)."""

    def save(self, state_dict, path: str):
        raise NotImplementedError

    def load(self, path: str, map_location=None):
        raise NotImplementedError

    def commit(self, info):
        """All files of checkpoint ``info`` are written; returns True when durable."""
        return True

    def makedirs(self, path, exist_ok=False):
        os.makedirs(path, exist_ok=exist_ok)

    def is_data_parallel_writer(self, dp_rank) -> bool:
        """Whether this data-parallel rank writes the replicated (model-state) files."""
        return dp_rank == 0

    def is_decoupled(self):
        return False

    def set_commit_info(self, info: CheckpointCommitInfo):
        """Engines that finish a checkpoint after ``save_checkpoint`` has returned remember what to commit."""

    def get_commit_info(self):
        return None

    def cleanup(self):
        pass
