"""Fused QKV projection sharding (reference: ``sharding/qkv.py:14-166``).

Layout here: ``wqkv [(hq + 2 hkv) * d, H]`` = all query heads, then all key heads, then all value heads.  A shard takes
its query-head range and the KV heads those query heads attend through (replicated when a KV group spans ranks) and
keeps the q|k|v order, so the local tensor is a valid fused QKV weight for ``(hq_local, hkv_local)``.
"""
from __future__ import annotations

from typing import Optional

import torch

from .attn import get_local_heads, head_ranges


def shard_qkv_param(param: Optional[torch.Tensor], shard_rank: int, num_shards: int, head_size: int,
                    n_heads_q: Optional[int] = None, n_heads_kv: Optional[int] = None) -> Optional[torch.Tensor]:
    if param is None:
        return None
    if num_shards == 1:
        return param
    if n_heads_q is None:
        if n_heads_kv is not None:
            raise ValueError("n_heads_kv needs n_heads_q")
        if param.shape[0] % (3 * head_size):
            raise ValueError("MHA sharding requires three equally sized q/k/v blocks of whole heads")
        n_heads_q = param.shape[0] // (3 * head_size)
    hkv = n_heads_q if n_heads_kv is None else n_heads_kv
    if param.shape[0] != (n_heads_q + 2 * hkv) * head_size:
        raise ValueError(f"QKV rows {param.shape[0]} != ({n_heads_q} + 2*{hkv}) * {head_size}")
    (q0, qn), (k0, kn) = head_ranges(shard_rank, num_shards, n_heads_q, n_heads_kv)
    d = head_size
    q_base, k_base, v_base = 0, n_heads_q * d, (n_heads_q + hkv) * d
    pieces = [param.narrow(0, q_base + q0 * d, qn * d), param.narrow(0, k_base + k0 * d, kn * d),
              param.narrow(0, v_base + k0 * d, kn * d)]
    return torch.cat(pieces, dim=0).contiguous()


def qkv_out_features(in_features: int, shard_rank: int, num_shards: int, head_size: int,
                     n_heads_q: Optional[int] = None, n_heads_kv: Optional[int] = None) -> int:
    if n_heads_q is None:
        n_heads_q = in_features // head_size
    q, kv = get_local_heads(shard_rank, num_shards, n_heads_q, n_heads_kv)
    return (q + 2 * kv) * head_size
