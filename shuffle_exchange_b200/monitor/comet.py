"""``CometMonitor`` under the reference's module path (``monitor/comet.py``); the implementation lives in ``monitor.py``."""
from .monitor import CometMonitor  # noqa: F401
