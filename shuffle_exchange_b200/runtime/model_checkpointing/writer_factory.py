"""Creates the file writer a checkpoint engine serialises into (reference
``runtime/model_checkpointing/writer_factory.py:14-95``): ``MOCK`` (counts bytes), ``PYTHON`` (buffered file) or ``FAST``
(double-buffered pinned staging + the native async-IO engine ``csrc/host/aio.cpp``; with ``use_gds`` the staging buffer
lives in device memory and is registered with the engine)."""
from __future__ import annotations

import torch

from ...io import FastFileWriter, FastFileWriterConfig, MockFileWriter, PyFileWriter
from ...utils.logging import logger
from ..swap_tensor.constants import (AIO_BLOCK_SIZE, AIO_INTRA_OP_PARALLELISM, AIO_OVERLAP_EVENTS, AIO_QUEUE_DEPTH,
                                     AIO_SINGLE_SUBMIT, AIO_USE_GDS)
from .constants import (CHECKPOINT_IO_BUFFER_DOUBLE, CHECKPOINT_IO_BUFFER_SIZE, CHECKPOINT_IO_MULTIPLIER,
                        CHECKPOINT_IO_STATISTICS, CHECKPOINT_WRITER_TYPE, CheckpointWriterType)


class CheckpointWriterFactory:

    def __init__(self, writer_config, aio_config, dp_writer_config):
        self._type = writer_config[CHECKPOINT_WRITER_TYPE]
        self._io_buffer_size = writer_config[CHECKPOINT_IO_BUFFER_SIZE]
        self._io_buffer_double = writer_config[CHECKPOINT_IO_BUFFER_DOUBLE]
        self._io_multiplier = writer_config[CHECKPOINT_IO_MULTIPLIER]
        self._data_parallel_writer = dp_writer_config
        self._show_statistics = bool(writer_config[CHECKPOINT_IO_STATISTICS]) and dp_writer_config is not None
        self._io_buffer = None
        self._dnvme_handle = None
        self._writer = None
        self._use_gds = False
        if self._type == CheckpointWriterType.FAST:
            self._use_gds = bool(aio_config.get(AIO_USE_GDS, False))
            (self._setup_for_gds if self._use_gds else self._setup_for_aio)(aio_config)
        logger.info(f"WriterFactory: type={self._type} data_parallel_writer={self._data_parallel_writer} "
                    f"show_statistics={self._show_statistics}")

    def create_writer(self, file_path, optimize_dp_state):
        assert self._writer is None, \
            f"Cannot create checkpoint writer for {file_path} because writer is currently used for " \
            f"{self._writer.file_path()}. Must call writer.release() before reusing to avoid this error."
        if self._type == CheckpointWriterType.MOCK:
            self._writer = MockFileWriter(file_path)
        elif self._type == CheckpointWriterType.PYTHON:
            self._writer = PyFileWriter(file_path)
        else:
            dp = self._data_parallel_writer
            if optimize_dp_state:
                n_writers, w_rank = dp.world_size * self._io_multiplier, dp.rank
                file_path = f"{file_path}-{w_rank}.{n_writers}"
            else:
                n_writers, w_rank = 1, 0
            cfg = FastFileWriterConfig(dnvme_handle=self._dnvme_handle, pinned_tensor=self._io_buffer,
                                       double_buffer=self._io_buffer_double, num_parallel_writers=n_writers,
                                       writer_rank=w_rank, global_rank=dp.global_rank if dp is not None else 0)
            self._writer = FastFileWriter(file_path=file_path, config=cfg)
        return self._writer

    def release_writer(self):
        self._writer.close()
        if self._show_statistics:
            self._writer._dump_state()
        self._writer = None

    def _new_handle(self, aio_config, gds: bool):
        from ...ops.aio import aio_handle
        kw = dict(block_size=aio_config[AIO_BLOCK_SIZE], queue_depth=aio_config[AIO_QUEUE_DEPTH],
                  single_submit=aio_config[AIO_SINGLE_SUBMIT], overlap_events=aio_config[AIO_OVERLAP_EVENTS],
                  intra_op_parallelism=aio_config[AIO_INTRA_OP_PARALLELISM])
        if gds:
            from ...ops.gds import gds_handle
            return gds_handle(**kw)
        return aio_handle(**kw)

    def _setup_for_aio(self, aio_config):
        self._io_buffer = torch.zeros(self._io_buffer_size, dtype=torch.uint8, device="cpu")
        if torch.cuda.is_available():
            self._io_buffer = self._io_buffer.pin_memory()
        self._dnvme_handle = self._new_handle(aio_config, gds=False)

    def _setup_for_gds(self, aio_config):
        dev = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else "cpu"
        self._io_buffer = torch.zeros(self._io_buffer_size, dtype=torch.uint8, device=dev)
        self._dnvme_handle = self._new_handle(aio_config, gds=True)
        self._dnvme_handle.pin_device_tensor(self._io_buffer)
