"""Monitor configuration under the reference's names (reference: ``monitor/config.py``)."""
from ..runtime.config import CometConfig, CSVConfig, MonitorConfig, TensorBoardConfig, WandbConfig  # noqa: F401

DeepSpeedMonitorConfig = MonitorConfig


def get_monitor_config(param_dict) -> MonitorConfig:
    return MonitorConfig(**{k: param_dict[k] for k in ("tensorboard", "wandb", "csv_monitor", "comet") if k in param_dict})
