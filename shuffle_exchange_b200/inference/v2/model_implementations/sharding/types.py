"""Which axis of a parameter a tensor-parallel shard cuts (reference: ``sharding/types.py:11-19``)."""
from enum import Enum


class ShardingType(Enum):
    OUTER_DIMENSION = 0     # rows of ``[out, in]`` — column-parallel linear; biases are cut too
    INNER_DIMENSION = 1     # columns of ``[out, in]`` — row-parallel linear; the bias is kept on rank 0 only


DEFAULT_SHARD_GRANULARITY = 32
