"""Checkpoint engine that serialises through :class:`~shuffle_exchange_b200.io.FastFileWriter` (pinned, double-buffered,
asynchronous NVMe writes) — reference: ``runtime/checkpoint_engine/fast_checkpoint_engine.py:17-60`` +
``runtime/model_checkpointing/writer_factory.py:14-90``."""
from __future__ import annotations

import torch

from ...io import FastFileWriter, FastFileWriterConfig, MockFileWriter, PyFileWriter
from ...ops.aio import aio_handle
from ...utils.logging import log_dist
from .checkpoint_engine import CheckpointEngine


class FastCheckpointEngine(CheckpointEngine):
    def __init__(self, config_params, dp_writer_config=None, optimize_dp_state=True):
        super().__init__(config_params)
        wc = getattr(config_params, "checkpoint_writer_config", None)
        self.writer_type = (getattr(wc, "type", "fast") or "fast").lower()
        self.show_statistics = bool(getattr(wc, "show_statistics", False))
        self.double_buffer = bool(getattr(wc, "io_buffer_double", True))
        nbytes = int(getattr(wc, "io_buffer_size", 1 << 28) or (1 << 28))
        self.optimize_dp_state = optimize_dp_state
        aio = getattr(config_params, "aio_config", None) or {}
        self._handle = None
        self._pinned = None
        if self.writer_type == "fast":
            self._handle = aio_handle(aio.get("block_size", 1 << 20), aio.get("queue_depth", 32),
                                      aio.get("single_submit", False), aio.get("overlap_events", True),
                                      max(1, aio.get("intra_op_parallelism", 4)))
            self._pinned = torch.empty(nbytes, dtype=torch.uint8, pin_memory=torch.cuda.is_available())

    def create(self, info):
        log_dist(f"[Fast] Checkpoint {info} is about to be saved!", ranks=[0])

    def _writer(self, path):
        if self.writer_type == "mock":
            return MockFileWriter(path)
        if self.writer_type == "python":
            return PyFileWriter(path)
        return FastFileWriter(path, FastFileWriterConfig(dnvme_handle=self._handle, pinned_tensor=self._pinned,
                                                         double_buffer=self.double_buffer))

    def save(self, state_dict, path: str):
        w = self._writer(path)
        torch.save(state_dict, w)
        w.close()
        if self.show_statistics:
            w._dump_state()

    def load(self, path: str, map_location=None):
        return torch.load(path, map_location=map_location, weights_only=False)

    def commit(self, info):
        log_dist(f"[Fast] Checkpoint {info} is ready now!", ranks=[0])
        return True
