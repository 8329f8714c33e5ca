"""``WandbMonitor`` under the reference's module path (``monitor/wandb.py``); the implementation lives in ``monitor.py``."""
from .monitor import WandbMonitor  # noqa: F401
