"""Head bookkeeping for tensor-parallel attention (reference: ``sharding/attn.py:9-62``).

MHA: query heads are dealt out as evenly as possible (leading ranks take the remainder).  GQA: either whole KV groups
are distributed (``n_heads_kv >= num_shards``; must divide) or one KV head is replicated over ``num_shards /
n_heads_kv`` ranks that split its query heads.
"""
from __future__ import annotations

from typing import Optional, Tuple


def _split(n: int, parts: int, idx: int) -> Tuple[int, int]:
    """(start, count) of part ``idx`` when ``n`` items are dealt to ``parts`` bins, larger bins first."""
    q, r = divmod(n, parts)
    return idx * q + min(idx, r), q + (1 if idx < r else 0)


def get_local_heads(shard_rank: int, num_shards: int, n_heads_q: int,
                    n_heads_kv: Optional[int] = None) -> Tuple[int, int]:
    q, kv = head_ranges(shard_rank, num_shards, n_heads_q, n_heads_kv)
    return q[1], kv[1]


def head_ranges(shard_rank: int, num_shards: int, n_heads_q: int,
                n_heads_kv: Optional[int] = None) -> Tuple[Tuple[int, int], Tuple[int, int]]:
    """((q_start, q_count), (kv_start, kv_count)) owned by ``shard_rank``."""
    if n_heads_q < num_shards:
        raise ValueError("There must be at least as many attention heads as there are shards.")
    if n_heads_kv is None or n_heads_kv == n_heads_q:
        r = _split(n_heads_q, num_shards, shard_rank)
        return r, r
    if n_heads_q % n_heads_kv:
        raise ValueError("Must be an even ratio between query and key/value heads.")
    ratio = n_heads_q // n_heads_kv
    if n_heads_kv >= num_shards:
        if n_heads_kv % num_shards:
            raise ValueError("If parallelizing groups, must be able to evenly distribute them.")
        per = n_heads_kv // num_shards
        return (shard_rank * per * ratio, per * ratio), (shard_rank * per, per)
    if num_shards % n_heads_kv:
        raise ValueError("If splitting a group across multiple shards, we must be able to distribute the groups evenly.")
    ranks_per_group = num_shards // n_heads_kv
    group, within = divmod(shard_rank, ranks_per_group)
    start, count = _split(ratio, ranks_per_group, within)
    return (group * ratio + start, count), (group, 1)
