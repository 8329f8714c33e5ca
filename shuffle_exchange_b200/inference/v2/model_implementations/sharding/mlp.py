"""MLP sharding (reference: ``sharding/mlp.py:14-77``): the first projection is column-parallel (gated activations
stack gate|up, each half cut separately), the second row-parallel.  ``is_moe`` marks tensors with a leading expert
dimension (``[E, out, in]`` weights, ``[E, out]`` biases)."""
from __future__ import annotations

from typing import Optional

import torch

from .types import DEFAULT_SHARD_GRANULARITY, ShardingType
from .utils import get_shard_endpoints, shard_param


def shard_mlp_1_param(param: Optional[torch.Tensor], shard_rank: int, num_shards: int, gated: bool = False,
                      is_moe: bool = False) -> Optional[torch.Tensor]:
    return shard_param(param, ShardingType.OUTER_DIMENSION, shard_rank, num_shards,
                       num_concatenated_matrices=2 if gated else 1, bias_dims=2 if is_moe else 1)


def shard_mlp_2_param(param: Optional[torch.Tensor], shard_rank: int, num_shards: int,
                      is_moe: bool = False) -> Optional[torch.Tensor]:
    return shard_param(param, ShardingType.INNER_DIMENSION, shard_rank, num_shards, bias_dims=2 if is_moe else 1)


def sharded_intermediate_dim(intermediate_size: int, num_shards: int, shard_rank: int) -> int:
    lo, hi = get_shard_endpoints(intermediate_size, shard_rank, num_shards, DEFAULT_SHARD_GRANULARITY)
    return hi - lo
