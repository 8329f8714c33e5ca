"""``csvMonitor`` under the reference's module path (``monitor/csv_monitor.py``); the implementation lives in ``monitor.py``."""
from .monitor import csvMonitor  # noqa: F401
