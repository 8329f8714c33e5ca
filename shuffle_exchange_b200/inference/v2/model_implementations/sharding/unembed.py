"""Unembedding sharding (reference: ``sharding/unembed.py:12-42``): ``[vocab, H]`` is cut along the vocabulary with
granularity 1 (vocabularies are rarely multiples of 32); ranks compute their logit columns and all-gather them."""
from __future__ import annotations

from typing import Optional

import torch

from .types import ShardingType
from .utils import get_shard_endpoints, shard_param


def shard_unembed_param(param: Optional[torch.Tensor], shard_rank: int, num_shards: int) -> Optional[torch.Tensor]:
    return shard_param(param, ShardingType.OUTER_DIMENSION, shard_rank, num_shards, granularity=1)


def sharded_unembed_dim(vocab_size: int, shard_rank: int, num_shards: int) -> int:
    lo, hi = get_shard_endpoints(vocab_size, shard_rank, num_shards, granularity=1)
    return hi - lo
