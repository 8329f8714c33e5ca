"""Topology summary handed to the writer factory (reference ``runtime/model_checkpointing/utils.py:14-84``)."""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import List, Optional

from ... import comm as dist
from .data_parallel_writer_factory import DataParallelWriterFactory

SOCKETS_PER_MACHINE = 2
LOCAL_RANK, CROSS_RANK, CROSS_SIZE = "LOCAL_RANK", "CROSS_RANK", "CROSS_SIZE"


@dataclass
class MPUInfo:
    pp_world_size: int
    pp_rank: int
    tp_world_size: int
    tp_rank: int
    dp_world_size: int
    dp_peer_ranks: List[int]
    dp_rank: int


@dataclass
class ExpertParallelInfo:
    ep_world_size: int
    ep_rank: int
    dp_world_size: int
    dp_peer_ranks: List[int]
    dp_rank: int


@dataclass
class UniversalParallelInfo:
    global_world_size: int
    global_rank: int
    local_rank: int
    mpu_info: Optional[MPUInfo]
    ep_info: Optional[ExpertParallelInfo]
    pure_dp: bool
    num_machines: int
    machine_rank: int
    num_sockets: int


def _create_model_parallel_info(mpu) -> MPUInfo:
    return MPUInfo(pp_world_size=mpu.get_pipeline_model_parallel_world_size(),
                   pp_rank=mpu.get_pipeline_model_parallel_rank(),
                   tp_world_size=mpu.get_tensor_model_parallel_world_size(),
                   tp_rank=mpu.get_tensor_model_parallel_rank(),
                   dp_world_size=mpu.get_data_parallel_world_size(),
                   dp_peer_ranks=mpu.get_data_parallel_group_ranks(),
                   dp_rank=mpu.get_data_parallel_rank())


def _create_expert_parallel_info(groups) -> ExpertParallelInfo:
    name = groups._get_max_expert_size_name()
    return ExpertParallelInfo(ep_world_size=groups._get_expert_parallel_world_size(name),
                              ep_rank=groups._get_expert_parallel_rank(name),
                              dp_world_size=groups._get_expert_data_parallel_world_size(name),
                              dp_peer_ranks=groups._get_expert_data_parallel_group_ranks(name),
                              dp_rank=groups._get_expert_data_parallel_rank(name))


def create_universal_parallel_info(groups, has_moe_layers) -> UniversalParallelInfo:
    world = dist.get_world_size() if dist.is_initialized() else 1
    rank = dist.get_rank() if dist.is_initialized() else 0
    mpu = getattr(groups, "mpu", None)
    n_machines = int(os.environ.get(CROSS_SIZE, 1))
    return UniversalParallelInfo(global_world_size=world, global_rank=rank,
                                 local_rank=int(os.environ.get(LOCAL_RANK, 0)),
                                 mpu_info=None if mpu is None else _create_model_parallel_info(mpu),
                                 ep_info=_create_expert_parallel_info(groups) if has_moe_layers else None,
                                 pure_dp=mpu is None or mpu.get_data_parallel_world_size() == world,
                                 num_machines=n_machines, machine_rank=int(os.environ.get(CROSS_RANK, 0)),
                                 num_sockets=n_machines * SOCKETS_PER_MACHINE)


def create_data_parallel_writer_config(groups, parallel_unit, zero_stage, has_moe_layers):
    info = create_universal_parallel_info(groups, has_moe_layers)
    return DataParallelWriterFactory(info, parallel_unit).create_config(zero_stage, has_moe_layers)
