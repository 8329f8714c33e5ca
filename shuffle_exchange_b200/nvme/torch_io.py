"""Engine module kept at the reference's path (``nvme/torch_io.py``); the implementation lives in :mod:`.io_engine`."""
from .io_engine import TorchIOEngine as Engine  # noqa: F401
from .io_engine import io_engine_multiprocessing  # noqa: F401
