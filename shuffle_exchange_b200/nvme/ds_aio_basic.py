"""Engine module kept at the reference's path (``nvme/ds_aio_basic.py``); the implementation lives in :mod:`.io_engine`."""
from .io_engine import AioBasicEngine as Engine  # noqa: F401
from .io_engine import io_engine_multiprocessing  # noqa: F401
