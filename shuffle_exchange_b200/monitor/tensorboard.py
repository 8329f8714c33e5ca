"""``TensorBoardMonitor`` under the reference's module path (``monitor/tensorboard.py``); the implementation lives in ``monitor.py``."""
from .monitor import TensorBoardMonitor  # noqa: F401
