"""Attention output projection sharding (reference: ``sharding/attn_out.py:14-112``): ``wo [H, hq * d]`` is cut along
its input columns to the rank's query heads (row-parallel); the bias stays on rank 0 so the all-reduce adds it once."""
from __future__ import annotations

from typing import Optional

import torch

from .attn import get_local_heads, head_ranges


def shard_attn_out_param(param: Optional[torch.Tensor], shard_rank: int, num_shards: int, head_size: int,
                         n_heads_q: Optional[int] = None, n_heads_kv: Optional[int] = None) -> Optional[torch.Tensor]:
    if param is None:
        return None
    if num_shards == 1:
        return param
    if param.dim() == 1:
        return param if shard_rank == 0 else None
    if n_heads_q is None:
        n_heads_q = param.shape[-1] // head_size
    (q0, qn), _ = head_ranges(shard_rank, num_shards, n_heads_q, n_heads_kv)
    return param[..., q0 * head_size:(q0 + qn) * head_size].contiguous()


def attn_out_in_features(out_features: int, shard_rank: int, num_shards: int, head_size: int,
                         n_heads_q: Optional[int] = None, n_heads_kv: Optional[int] = None) -> int:
    if n_heads_q is None:
        n_heads_q = out_features // head_size
    return get_local_heads(shard_rank, num_shards, n_heads_q, n_heads_kv)[0] * head_size
