"""``tensor_parallel`` config block (reference: ``runtime/tensor_parallel/config.py:11-81``)."""
from __future__ import annotations

from enum import Enum
from typing import Any, Optional

import torch
from pydantic import Field, model_validator

from ..config_utils import DeepSpeedConfigModel


class AUTOTP_MODE(Enum):
    TRAINING = "TRAINING"
    INFERENCE = "INFERENCE"


class TPConfig(DeepSpeedConfigModel):
    """``tp_size`` ranks per tensor-parallel group; ``tp_grain_size`` = granularity of uneven column splits (set to the
    GEMM tile multiple you want every shard to keep); ``mpu`` / ``tp_group`` = externally created groups."""
    tp_size: int = 1
    tp_grain_size: int = 1
    mpu: Any = None
    tp_group: Any = None


class TPTrainingConfig(DeepSpeedConfigModel):
    dtype: Any = torch.float16
    autotp_size: int = 0
    tp_overlap_comm: bool = False
    tensor_parallel: TPConfig = Field({}, alias="tp")
    injection_policy_tuple: Optional[tuple] = None
    keep_module_on_host: bool = False
    replace_with_kernel_inject: bool = Field(False, alias="kernel_inject")

    @model_validator(mode="after")
    def _sync_sizes(self):
        if self.autotp_size and self.tensor_parallel.tp_size == 1:
            self.tensor_parallel.tp_size = self.autotp_size
        return self


def get_tensor_parallel_config(ds_config: dict) -> TPTrainingConfig:
    if "tensor_parallel" in ds_config:
        return TPTrainingConfig(**ds_config["tensor_parallel"])
    return TPTrainingConfig()
