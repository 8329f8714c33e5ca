"""Shard arithmetic shared by every v2 parameter kind (reference: ``sharding/utils.py:13-103``).

A dimension is divided in units of ``granularity`` elements; the first ``units % num_shards`` ranks take one unit more.
Boundaries are a pure function of ``(dim_size, rank, num_shards, granularity)`` so every rank derives every other
rank's slice without communication.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from .types import DEFAULT_SHARD_GRANULARITY, ShardingType


def get_shard_endpoints(dim_size: int, shard_rank: int, num_shards: int,
                        granularity: int = DEFAULT_SHARD_GRANULARITY) -> Tuple[int, int]:
    if dim_size % granularity:
        raise ValueError(f"dimension {dim_size} is not a multiple of the shard granularity {granularity}")
    if not 0 <= shard_rank < num_shards:
        raise ValueError(f"shard rank {shard_rank} outside [0, {num_shards})")
    units = dim_size // granularity
    q, r = divmod(units, num_shards)
    lo = shard_rank * q + min(shard_rank, r)
    hi = lo + q + (1 if shard_rank < r else 0)
    return lo * granularity, hi * granularity


def shard_param(param: Optional[torch.Tensor], shard_mode: ShardingType, shard_rank: int, num_shards: int,
                num_concatenated_matrices: int = 1, granularity: int = DEFAULT_SHARD_GRANULARITY,
                bias_dims: int = 1) -> Optional[torch.Tensor]:
    """Cut ``param`` for ``shard_rank``.

    ``num_concatenated_matrices`` > 1 means dim ``-2`` (or the only dim of a bias) is several equally sized matrices
    stacked (gate|up): each is cut separately and the pieces re-stacked, so the local tensor keeps the same fused
    structure.  ``bias_dims`` is the rank at (or below) which the tensor is a bias: 1 for dense layers, 2 for
    per-expert biases ``[E, out]``.
    """
    if param is None:
        return None
    if num_shards == 1:
        return param
    is_bias = param.dim() <= bias_dims
    if shard_mode == ShardingType.INNER_DIMENSION:
        if is_bias:                                  # added once, after the all-reduce's owner
            return param if shard_rank == 0 else None
        lo, hi = get_shard_endpoints(param.shape[-1], shard_rank, num_shards, granularity)
        return param[..., lo:hi].contiguous()
    axis = param.dim() - 1 if is_bias else param.dim() - 2
    full = param.shape[axis]
    if full % num_concatenated_matrices:
        raise ValueError(f"dimension {full} does not hold {num_concatenated_matrices} stacked matrices")
    each = full // num_concatenated_matrices
    lo, hi = get_shard_endpoints(each, shard_rank, num_shards, granularity)
    pieces = [param.narrow(axis, m * each + lo, hi - lo) for m in range(num_concatenated_matrices)]
    return (pieces[0] if len(pieces) == 1 else torch.cat(pieces, dim=axis)).contiguous()
