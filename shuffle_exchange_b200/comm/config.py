"""``comms_logger`` config block (reference: deepspeed/comm/config.py, comm/constants.py:20-51)."""
from typing import List

from pydantic import BaseModel


class CommsLoggerConfig(BaseModel):
    enabled: bool = False
    prof_all: bool = True
    prof_ops: List[str] = []
    verbose: bool = False
    debug: bool = False


class DeepSpeedCommsConfig:
    def __init__(self, ds_config: dict):
        self.comms_logger_enabled = "comms_logger" in ds_config
        self.comms_logger = CommsLoggerConfig(**ds_config.get("comms_logger", {}))
