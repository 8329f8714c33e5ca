"""Paged KV cache (reference: ``inference/v2/ragged/kv_cache.py:40-208``): one tensor
``[n_layers, n_blocks, block_size, 2, n_kv_heads, head_dim]`` per cache group + a :class:`BlockedAllocator`.  Sizing
follows the reference's policy: either an explicit block count or "fill ``memory_fraction`` of free HBM"."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple

import torch

from .blocked_allocator import BlockedAllocator


@dataclass
class KVCacheConfig:
    block_size: int = 128
    num_allocation_groups: int = 1
    cache_shape: Tuple[int, int, int] = (1, 1, 1)        # (n_layers, n_kv_heads, head_dim)
    cache_dtype: torch.dtype = torch.bfloat16
    max_blocks_per_allocation_group: int = 64


def split_kv(kv_cache: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """Key and value views ``[n_blocks, block_size, n_kv_heads, head_dim]`` of one layer's cache
    ``[n_blocks, block_size, 2, n_kv_heads, head_dim]`` (reference ``kv_cache.py:22-37``)."""
    if kv_cache.ndim != 5 or kv_cache.shape[2] != 2:
        raise ValueError(f"expected a [blocks, block_size, 2, heads, head_dim] cache, got {tuple(kv_cache.shape)}")
    return kv_cache.select(2, 0), kv_cache.select(2, 1)


class BlockedKVCache:
    def __init__(self, config: KVCacheConfig, num_blocks: Optional[int] = None, memory_fraction: float = 0.8,
                 device=None, reserve_bytes: Optional[int] = None, mp_group=None) -> None:
        self._config = config
        n_layers, n_kv, d = config.cache_shape
        elem = torch.tensor([], dtype=config.cache_dtype).element_size()
        per_block = n_layers * config.block_size * 2 * n_kv * d * elem
        dev = device if device is not None else ("cuda" if torch.cuda.is_available() else "cpu")
        if num_blocks is None:
            if torch.device(dev).type == "cuda":
                free, _ = torch.cuda.mem_get_info()
                usable = free - reserve_bytes if reserve_bytes is not None else int(free * memory_fraction)
                num_blocks = max(1, int(usable) // per_block)
            else:
                num_blocks = 256
            if mp_group is not None:
                # tensor-parallel ranks must agree on the block count (reference ``kv_cache.py:116-124``: MIN all-reduce)
                import torch.distributed as dist
                t = torch.tensor([num_blocks], dtype=torch.int64, device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MIN, group=mp_group)
                num_blocks = int(t.item())
        self._num_blocks = num_blocks
        self._cache = torch.zeros(n_layers, num_blocks, config.block_size, 2, n_kv, d, dtype=config.cache_dtype, device=dev)
        self._allocator = BlockedAllocator(num_blocks)

    def reserve(self, num_blocks: int) -> torch.Tensor:
        return self._allocator.allocate(num_blocks)

    def free(self, blocks) -> None:
        self._allocator.free(blocks)

    def get_cache(self, layer: int) -> torch.Tensor:
        return self._cache[layer]

    @property
    def free_blocks(self) -> int:
        return self._allocator.free_blocks

    @property
    def num_blocks(self) -> int:
        return self._num_blocks

    @property
    def block_size(self) -> int:
        return self._config.block_size
