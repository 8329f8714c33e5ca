"""Embedding sharding (reference: ``sharding/embedding.py:12-34``): ``[vocab, H]`` is cut along the hidden dimension —
every rank looks up every token and holds ``H / tp`` of each row; an all-gather along the feature axis rebuilds the
residual stream."""
from __future__ import annotations

from typing import Optional

import torch

from .types import ShardingType
from .utils import get_shard_endpoints, shard_param


def shard_embedding_param(param: Optional[torch.Tensor], shard_rank: int, num_shards: int) -> Optional[torch.Tensor]:
    return shard_param(param, ShardingType.INNER_DIMENSION, shard_rank, num_shards)


def sharded_embedding_dim(embedding_size: int, shard_rank: int, num_shards: int) -> int:
    lo, hi = get_shard_endpoints(embedding_size, shard_rank, num_shards)
    return hi - lo
