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

    # ----------------------------------------------------------------------- host offload of blocks
    def offload(self, blocks) -> torch.Tensor:
        """Copy ``blocks`` (all layers) to host memory and return their block ids to the free list.  Returns the host
        tensor ``[n_layers, len(blocks), block_size, 2, n_kv_heads, head_dim]`` (pinned when the cache is on a GPU) —
        the caller keeps it and hands it back to :meth:`restore`.  (The reference declares both methods and raises
        ``NotImplementedError``, ``kv_cache.py:166-184``.)"""
        ids = torch.as_tensor(blocks, dtype=torch.long).reshape(-1)
        picked = self._cache.index_select(1, ids.to(self._cache.device))
        if self._cache.is_cuda:
            host = torch.empty(picked.shape, dtype=picked.dtype, device="cpu", pin_memory=True)
            host.copy_(picked, non_blocking=True)
            torch.cuda.current_stream(self._cache.device).synchronize()   # blocks are reusable once this returns
        else:
            host = picked.clone()
        self._allocator.free(ids.to(torch.int32))
        return host

    def restore(self, host: torch.Tensor) -> torch.Tensor:
        """Allocate as many blocks as ``host`` holds, copy it back and return the new block ids (they differ from the
        ids the data was offloaded from; the sequence's block table is what gives them meaning)."""
        n = host.shape[1]
        ids = self._allocator.allocate(n)
        self._cache.index_copy_(1, ids.to(device=self._cache.device, dtype=torch.long),
                                host.to(self._cache.device, non_blocking=True))
        return ids

    @property
    def num_caches(self) -> int:
        return self._cache.shape[0]

    @property
    def free_blocks(self) -> int:
        return self._allocator.free_blocks

    @property
    def num_blocks(self) -> int:
        return self._num_blocks

    @property
    def block_size(self) -> int:
        return self._config.block_size
