"""``zero_optimization`` config block.

Keys and defaults follow the reference (deepspeed/runtime/zero/config.py; listed in SURVEY.md §5.6)
plus one addition: ``replica_averaging`` carries the fork's four knobs (``method``, ``rings``,
``shuffle_step``, ``slice_count``) and an explicit ``seed`` so the shuffle schedule is reproducible;
the same values are still accepted as ``initialize()`` kwargs for drop-in compatibility.
"""
from __future__ import annotations

import sys
from enum import Enum
from typing import Dict, Optional

from pydantic import Field, model_validator

from ...utils.logging import logger
from ..config_utils import DeepSpeedConfigModel, get_scalar_param, pp_int
from .offload_config import (DeepSpeedZeroOffloadOptimizerConfig, DeepSpeedZeroOffloadParamConfig,
                             OffloadDeviceEnum)

ZERO_OPTIMIZATION = "zero_optimization"


class ZeroStageEnum(int, Enum):
    disabled = 0
    optimizer_states = 1
    gradients = 2
    weights = 3
    max_stage = 3


class ReplicaAveragingConfig(DeepSpeedConfigModel):
    """Decentralised parameter averaging between ZeRO "slices" (the fork's feature)."""
    method: str = "RR"          # RR | shuffle | H-RR | Gossip
    rings: int = Field(8, ge=1)
    shuffle_step: int = Field(50, ge=1)
    slice_count: int = Field(2, ge=1)
    seed: int = 1234
    fused: Optional[bool] = None   # None = use the symmetric-memory kernel when available


def read_zero_config_deprecated(param_dict):
    zero_config_dict = {"stage": 1 if param_dict[ZERO_OPTIMIZATION] else 0}
    if zero_config_dict["stage"] > 0:
        zero_config_dict["allgather_bucket_size"] = get_scalar_param(param_dict, "allgather_size", 5e8)
    logger.warning("DeepSpeedConfig: this format of ZeRO optimization setup is deprecated. "
                   f"Please use the following format: {ZERO_OPTIMIZATION}: {{'stage': [0|1|2|3], ...}}")
    return zero_config_dict


def get_zero_config(param_dict) -> "DeepSpeedZeroConfig":
    if ZERO_OPTIMIZATION in param_dict:
        zero_config_dict = param_dict[ZERO_OPTIMIZATION]
        if isinstance(zero_config_dict, bool):
            zero_config_dict = read_zero_config_deprecated(param_dict)
    else:
        zero_config_dict = {}
    return DeepSpeedZeroConfig(**zero_config_dict)


class DeepSpeedZeroConfig(DeepSpeedConfigModel):
    stage: ZeroStageEnum = 0
    contiguous_gradients: bool = True
    reduce_scatter: bool = True
    reduce_bucket_size: int = Field(pp_int(5e8), ge=0)
    use_multi_rank_bucket_allreduce: bool = True
    allgather_partitions: bool = True
    allgather_bucket_size: int = Field(pp_int(5e8), ge=0)
    overlap_comm: Optional[bool] = None      # None -> True for stage 3, False otherwise
    load_from_fp32_weights: bool = True
    elastic_checkpoint: bool = False

    offload_param: Optional[DeepSpeedZeroOffloadParamConfig] = None
    offload_optimizer: Optional[DeepSpeedZeroOffloadOptimizerConfig] = None

    sub_group_size: int = Field(pp_int(1e9), ge=0)
    cpu_offload_param: Optional[bool] = Field(
        None, json_schema_extra={"deprecated": True, "new_param": "offload_param",
                                 "new_param_fn": (lambda val: DeepSpeedZeroOffloadParamConfig(device=OffloadDeviceEnum.cpu) if val else None)})
    cpu_offload_use_pin_memory: Optional[bool] = Field(
        None, json_schema_extra={"deprecated": True, "new_param": "offload_param or offload_optimizer",
                                 "set_new_param": False})
    cpu_offload: Optional[bool] = Field(
        None, json_schema_extra={"deprecated": True, "new_param": "offload_optimizer",
                                 "new_param_fn": (lambda val: DeepSpeedZeroOffloadOptimizerConfig(device=OffloadDeviceEnum.cpu) if val else None)})

    prefetch_bucket_size: int = Field(pp_int(5e7), ge=0, alias="stage3_prefetch_bucket_size")
    param_persistence_threshold: int = Field(pp_int(1e5), ge=0, alias="stage3_param_persistence_threshold")
    model_persistence_threshold: int = Field(pp_int(sys.maxsize, "sys.maxsize"), ge=0,
                                             alias="stage3_model_persistence_threshold")
    max_live_parameters: int = Field(pp_int(1e9), ge=0, alias="stage3_max_live_parameters")
    max_reuse_distance: int = Field(pp_int(1e9), ge=0, alias="stage3_max_reuse_distance")
    gather_16bit_weights_on_model_save: bool = Field(False, alias="stage3_gather_16bit_weights_on_model_save")
    module_granularity_threshold: int = Field(pp_int(0), alias="stage3_module_granularity_threshold")
    use_all_reduce_for_fetch_params: bool = Field(False, alias="stage3_use_all_reduce_for_fetch_params")
    stage3_gather_fp16_weights_on_model_save: bool = Field(
        False, json_schema_extra={"deprecated": True, "new_param": "gather_16bit_weights_on_model_save"})

    ignore_unused_parameters: bool = True
    legacy_stage1: bool = False
    round_robin_gradients: bool = False
    zero_hpz_partition_size: int = Field(1, ge=0)
    zero_quantized_weights: bool = False
    zero_quantized_nontrainable_weights: bool = False
    zero_quantized_gradients: bool = False
    zeropp_loco_param: Optional[Dict[str, float]] = None
    mics_shard_size: int = Field(-1, json_schema_extra={"new_param": "mics_shard_size"})
    mics_hierarchical_params_gather: bool = False
    memory_efficient_linear: bool = True
    pipeline_loading_checkpoint: bool = False
    override_module_apply: bool = True
    log_trace_cache_warnings: bool = False

    # ---- additions of this framework -------------------------------------------------------
    replica_averaging: Optional[ReplicaAveragingConfig] = None
    fused_collectives: Optional[bool] = None   # None = auto (symmetric heap present and world > 1)
    fused_step: bool = True                    # run Adam on a unit's shard as soon as its grads are reduced

    @model_validator(mode="after")
    def overlap_comm_valid(self):
        if self.overlap_comm is None:
            object.__setattr__(self, "overlap_comm", self.stage == ZeroStageEnum.weights)
        return self

    @model_validator(mode="after")
    def offload_ratio_check(self):
        oc = self.offload_optimizer
        if oc is not None and oc.ratio < 1.0:
            assert self.stage == ZeroStageEnum.weights, "Partial offloading only supported for ZeRO Stage 3."
        return self
