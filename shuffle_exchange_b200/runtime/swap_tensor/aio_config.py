"""Parse the ``aio`` section of the config (reference: ``runtime/swap_tensor/aio_config.py:20-36``)."""
from .constants import *  # noqa: F401,F403

AIO_DEFAULT_DICT = {
    AIO_BLOCK_SIZE: AIO_BLOCK_SIZE_DEFAULT, AIO_QUEUE_DEPTH: AIO_QUEUE_DEPTH_DEFAULT,
    AIO_INTRA_OP_PARALLELISM: AIO_INTRA_OP_PARALLELISM_DEFAULT, AIO_SINGLE_SUBMIT: AIO_SINGLE_SUBMIT_DEFAULT,
    AIO_OVERLAP_EVENTS: AIO_OVERLAP_EVENTS_DEFAULT, AIO_USE_GDS: AIO_USE_GDS_DEFAULT,
}


def get_aio_config(param_dict):
    section = param_dict.get(AIO) if isinstance(param_dict, dict) else None
    if not section:
        return dict(AIO_DEFAULT_DICT)
    out = dict(AIO_DEFAULT_DICT)
    out.update({k: section[k] for k in AIO_DEFAULT_DICT if k in section})
    if AIO_THREAD_COUNT in section and AIO_INTRA_OP_PARALLELISM not in section:      # older name
        out[AIO_INTRA_OP_PARALLELISM] = section[AIO_THREAD_COUNT]
    return out
