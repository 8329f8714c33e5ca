"""Process-group construction for MiCS (reference: ``runtime/zero/mics_utils.py:20-203``).

The data-parallel world of size ``W`` is cut into ``W / shard_size`` *shard groups* of consecutive ranks (ZeRO-3
partitioning happens inside one) and ``shard_size`` *replicate groups* (ranks holding the same shard index).  With
``hierarchical_allgather`` on more than one node, a shard group is further split into an inter-node group (same
local index on every node of the shard group) and an intra-node group.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional

import numpy as np

from ... import comm as dist
from ...utils.logging import logger


@dataclass
class MiCS_CommGroups:
    param_shard_group: object = None
    param_shard_size: int = -1
    param_shard_rank: int = -1
    param_repli_group: object = None
    param_repli_size: int = -1
    param_repli_rank: int = -1
    param_intra_node_group: object = None
    param_inter_node_shard_group: object = None


def _generate_mics_config(world_size: int, ndev_per_node: int, shard_size: int, pp_size: int = 1):
    """Rank lists: ``shard_groups`` (consecutive blocks), ``replicate_groups`` (strided), optional hierarchy."""
    assert world_size % pp_size == 0
    assert (world_size // pp_size) % shard_size == 0, \
        f"dp group size is not dividable by dp_shard_size, (world_size {world_size}, pp_size {pp_size}, dp_shard_size {shard_size})"
    config = {"shard_groups": [], "replicate_groups": [], "span_nodes": 1}
    ranks = np.arange(world_size).reshape(pp_size, -1)
    for row in ranks:
        blocks = row.reshape(-1, shard_size)
        config["shard_groups"] += blocks.tolist()
        config["replicate_groups"] += blocks.T.tolist()
    if shard_size > ndev_per_node:
        assert shard_size % ndev_per_node == 0, "partition size must be a multiple of the devices per node"
        config["span_nodes"] = shard_size // ndev_per_node
    return config


def _sizes_all_same(groups: List[List[int]]) -> bool:
    return len({len(g) for g in groups}) <= 1


def create_mics_comm_groups(shard_size: int, dp_group, hierarchical_allgather: bool = False, mpu=None) -> MiCS_CommGroups:
    import torch
    ndev = max(1, torch.cuda.device_count()) if torch.cuda.is_available() else dist.get_world_size()
    world, rank = dist.get_world_size(), dist.get_rank()
    if mpu is not None:
        assert dp_group == mpu.get_data_parallel_group()
    cfg = _generate_mics_config(world, ndev, shard_size, 1)
    groups = MiCS_CommGroups()
    assert _sizes_all_same(cfg["shard_groups"]) and _sizes_all_same(cfg["replicate_groups"])
    for ranks in cfg["shard_groups"]:
        g = dist.new_group(ranks)
        if rank in ranks:
            groups.param_shard_group, groups.param_shard_size = g, len(ranks)
            groups.param_shard_rank = ranks.index(rank)
            logger.info(f"rank {rank}, shard group {ranks.index(rank)}/{len(ranks)}")
    for ranks in cfg["replicate_groups"]:
        if len(ranks) > 1:
            g = dist.new_group(ranks)
            if rank in ranks:
                groups.param_repli_group, groups.param_repli_size = g, len(ranks)
                groups.param_repli_rank = ranks.index(rank)
        elif rank in ranks:
            groups.param_repli_group, groups.param_repli_size, groups.param_repli_rank = None, 1, 0
    span = cfg["span_nodes"]
    if hierarchical_allgather and span > 1:
        for ranks in cfg["shard_groups"]:
            grid = np.array(ranks).reshape(span, ndev)
            for local in range(ndev):                       # same local index across the nodes of the shard group
                inter = grid[:, local].tolist()
                g = dist.new_group(inter)
                if rank in inter:
                    groups.param_inter_node_shard_group = g
            for node in range(span):
                intra = grid[node].tolist()
                g = dist.new_group(intra)
                if rank in intra:
                    groups.param_intra_node_group = g
    return groups
