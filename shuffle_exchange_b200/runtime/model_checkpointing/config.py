"""Dict-style accessors for the ``checkpoint`` section (reference ``runtime/model_checkpointing/config.py:9-77``); the
engine's pydantic ``CheckpointConfig`` (``runtime/config.py``) parses the same keys — these helpers serve code that
works on the raw dict (writer factories, the decoupled checkpoint engine)."""
from .constants import *  # noqa: F401,F403
from .constants import (CHECKPOINT, CHECKPOINT_DATA_PARALLEL, CHECKPOINT_DATA_PARALLEL_DEFAULT,
                        CHECKPOINT_DATA_PARALLEL_UNITS, CHECKPOINT_IO_BUFFER_DOUBLE,
                        CHECKPOINT_IO_BUFFER_DOUBLE_DEFAULT, CHECKPOINT_IO_BUFFER_SIZE,
                        CHECKPOINT_IO_BUFFER_SIZE_DEFAULT, CHECKPOINT_IO_MULTIPLIER, CHECKPOINT_IO_MULTIPLIER_DEFAULT,
                        CHECKPOINT_IO_STATISTICS, CHECKPOINT_IO_STATISTICS_DEFAULT, CHECKPOINT_SERIALIZATION,
                        CHECKPOINT_SERIALIZATION_DEFAULT, CHECKPOINT_TAG_VALIDATION, CHECKPOINT_TAG_VALIDATION_DEFAULT,
                        CHECKPOINT_TAG_VALIDATION_MODES, CHECKPOINT_WRITER, CHECKPOINT_WRITER_DECOUPLED,
                        CHECKPOINT_WRITER_DECOUPLED_DEFAULT, CHECKPOINT_WRITER_DEFAULT, CHECKPOINT_WRITER_TYPE,
                        CHECKPOINT_WRITER_TYPE_DEFAULT, CHECKPOINT_WRITER_TYPES)

VALID_VALUES = {
    CHECKPOINT_TAG_VALIDATION: CHECKPOINT_TAG_VALIDATION_MODES,
    CHECKPOINT_WRITER_TYPE: CHECKPOINT_WRITER_TYPES,
    CHECKPOINT_DATA_PARALLEL: CHECKPOINT_DATA_PARALLEL_UNITS,
}
CHECKPOINT_DEFAULT_DICT = {
    CHECKPOINT_TAG_VALIDATION: CHECKPOINT_TAG_VALIDATION_DEFAULT,
    CHECKPOINT_SERIALIZATION: CHECKPOINT_SERIALIZATION_DEFAULT,
    CHECKPOINT_WRITER: CHECKPOINT_WRITER_DEFAULT,
}


def _validate_config_values(config_name, config_dict, valid_values):
    for key, value in config_dict.items():
        if value is not None and key in valid_values:
            assert value in valid_values[key], \
                f"{config_name} contains invalid value {value} for {key}, expecting one of {valid_values[key]}"


def _upper(value):
    return value if value is None else str(value).upper()


def get_checkpoint_writer_config(param_dict):
    writer = param_dict.get(CHECKPOINT_WRITER, None)
    if writer is None:
        return CHECKPOINT_WRITER_DEFAULT
    cfg = {
        CHECKPOINT_WRITER_TYPE: _upper(writer.get(CHECKPOINT_WRITER_TYPE, CHECKPOINT_WRITER_TYPE_DEFAULT)),
        CHECKPOINT_IO_BUFFER_SIZE: int(writer.get(CHECKPOINT_IO_BUFFER_SIZE, CHECKPOINT_IO_BUFFER_SIZE_DEFAULT)),
        CHECKPOINT_IO_BUFFER_DOUBLE: writer.get(CHECKPOINT_IO_BUFFER_DOUBLE, CHECKPOINT_IO_BUFFER_DOUBLE_DEFAULT),
        CHECKPOINT_IO_STATISTICS: writer.get(CHECKPOINT_IO_STATISTICS, CHECKPOINT_IO_STATISTICS_DEFAULT),
        CHECKPOINT_DATA_PARALLEL: _upper(writer.get(CHECKPOINT_DATA_PARALLEL, CHECKPOINT_DATA_PARALLEL_DEFAULT)),
        CHECKPOINT_WRITER_DECOUPLED: writer.get(CHECKPOINT_WRITER_DECOUPLED, CHECKPOINT_WRITER_DECOUPLED_DEFAULT),
        CHECKPOINT_IO_MULTIPLIER: int(writer.get(CHECKPOINT_IO_MULTIPLIER, CHECKPOINT_IO_MULTIPLIER_DEFAULT)),
    }
    _validate_config_values(CHECKPOINT_WRITER, cfg, VALID_VALUES)
    return cfg


def get_checkpoint_config(param_dict):
    section = param_dict.get(CHECKPOINT, None)
    if section is None:
        return dict(CHECKPOINT_DEFAULT_DICT)
    cfg = {
        CHECKPOINT_TAG_VALIDATION: str(section.get(CHECKPOINT_TAG_VALIDATION,
                                                   CHECKPOINT_TAG_VALIDATION_DEFAULT)).upper(),
        CHECKPOINT_SERIALIZATION: section.get(CHECKPOINT_SERIALIZATION, CHECKPOINT_SERIALIZATION_DEFAULT),
        CHECKPOINT_WRITER: get_checkpoint_writer_config(section),
    }
    _validate_config_values(CHECKPOINT, cfg, VALID_VALUES)
    return cfg
