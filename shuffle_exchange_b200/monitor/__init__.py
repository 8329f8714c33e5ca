from .monitor import MonitorMaster, TensorBoardMonitor, WandbMonitor, csvMonitor, CometMonitor
