"""Engine module kept at the reference's path (``nvme/torch_fastio_engine.py``); the implementation lives in :mod:`.io_engine`."""
from .io_engine import TorchFastIOEngine as Engine  # noqa: F401
from .io_engine import io_engine_multiprocessing  # noqa: F401
