"""Free-list allocator of KV-cache block ids (reference: ``inference/v2/ragged/blocked_allocator.py:11-105``).
The free list is an intrusive singly linked list stored in one int32 tensor: ``_next[i]`` is the block that follows
``i`` on the free list; allocation pops ``n`` heads, ``free`` pushes blocks back.  O(n) per call, no Python lists."""
from __future__ import annotations

from typing import Iterable, Union

import torch


class BlockedAllocator:
    def __init__(self, num_blocks: int) -> None:
        if num_blocks < 1:
            raise ValueError(f"Blocked KV-cache must have at least 1 block, provided {num_blocks}")
        self._num_blocks = num_blocks
        self._next = torch.arange(1, num_blocks + 1, dtype=torch.int32, device="cpu")
        self._head = 0
        self._free_blocks = num_blocks

    def allocate(self, num_blocks: int) -> torch.Tensor:
        if num_blocks > self._free_blocks:
            raise ValueError(f"Not enough free blocks in the KV-cache to allocate {num_blocks} blocks")
        out = torch.zeros(num_blocks, dtype=torch.int32)
        for i in range(num_blocks):
            out[i] = self._head
            self._head = int(self._next[self._head])
            self._next[out[i]] = -1            # marks "in use" (double-free detection)
        self._free_blocks -= num_blocks
        return out

    def free(self, blocks: Union[Iterable[int], int]) -> None:
        if isinstance(blocks, int):
            blocks = [blocks]
        blocks = [int(b) for b in (blocks.tolist() if torch.is_tensor(blocks) else blocks)]
        for b in blocks:
            if b < 0 or b >= self._num_blocks:
                raise ValueError(f"Invalid block {b} provided to free")
            if self._next[b] != -1:
                raise ValueError(f"Block {b} is already free")
        for b in blocks:
            self._next[b] = self._head
            self._head = b
            self._free_blocks += 1

    @property
    def free_blocks(self) -> int:
        return self._free_blocks
