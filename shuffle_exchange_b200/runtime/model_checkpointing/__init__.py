"""Parallel checkpoint-writer selection (reference ``deepspeed/runtime/model_checkpointing``)."""
from .config import get_checkpoint_config, get_checkpoint_writer_config
from .constants import CheckpointDataParallel, CheckpointWriterType, ValidationMode
from .data_parallel_writer_factory import DataParallelWriterConfig, DataParallelWriterFactory
from .utils import (ExpertParallelInfo, MPUInfo, UniversalParallelInfo, create_data_parallel_writer_config,
                    create_universal_parallel_info)
from .writer_factory import CheckpointWriterFactory
