"""Checkpoint engine interface (reference: deepspeed/runtime/checkpoint_engine/checkpoint_engine.py:20-60)."""


class CheckpointEngine:
    def __init__(self, config_params=None):
        self.config_params = config_params

    def create(self, info):
        """Announce the start of a checkpoint identified by ``info`` (tag)."""

    def save(self, state_dict, path: str):
        raise NotImplementedError

    def load(self, path: str, map_location=None):
        raise NotImplementedError

    def commit(self, info):
        """All files of checkpoint ``info`` are written; returns True when durable."""
        return True

    def makedirs(self, path, exist_ok=False):
        import os
        os.makedirs(path, exist_ok=exist_ok)

    def is_decoupled(self):
        return False

    def cleanup(self):
        pass
