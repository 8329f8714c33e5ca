"""Per-sequence bookkeeping (reference: ``inference/v2/ragged/sequence_descriptor.py:58-280``): tokens already in the
cache, tokens in flight in the current forward, and the KV block ids the sequence owns."""
from __future__ import annotations

from typing import List

import torch


class BaseSequenceDescriptor:
    """What the scheduler-side checks read from a sequence (reference ``sequence_descriptor.py:15-33``)."""

    @property
    def seen_tokens(self) -> int:
        raise NotImplementedError

    @property
    def cur_allocated_blocks(self) -> int:
        raise NotImplementedError


class PlaceholderSequenceDescriptor(BaseSequenceDescriptor):
    """Stand-in for a sequence that is not tracked yet: lets ``query`` / ``can_schedule`` ask the model for the KV
    requirement of a new request without allocating a block table (reference ``sequence_descriptor.py:35-56``)."""

    def __init__(self, seen_tokens: int = 0, cur_allocated_blocks: int = 0, kv_blocks_ptr: int = 0) -> None:
        self._seen_tokens, self._cur_allocated_blocks, self._kv_blocks_ptr = seen_tokens, cur_allocated_blocks, kv_blocks_ptr

    @property
    def seen_tokens(self) -> int:
        return self._seen_tokens

    @property
    def cur_allocated_blocks(self) -> int:
        return self._cur_allocated_blocks

    @property
    def kv_blocks_ptr(self) -> int:
        return self._kv_blocks_ptr


class DSSequenceDescriptor(BaseSequenceDescriptor):
    def __init__(self, tracking_id: int, max_blocks: int) -> None:
        self._tracking_id = tracking_id
        self._seen_tokens = 0
        self._in_flight_tokens = 0
        self._blocks = torch.zeros(max_blocks, dtype=torch.int32)
        self._n_blocks = 0
        self._host_kv = None          # host copy of the KV blocks while the sequence is offloaded

    @property
    def tracking_id(self) -> int:
        return self._tracking_id

    @property
    def seen_tokens(self) -> int:
        return self._seen_tokens

    @property
    def in_flight_tokens(self) -> int:
        return self._in_flight_tokens

    @property
    def cur_allocated_blocks(self) -> int:
        return self._n_blocks

    @property
    def max_context(self) -> int:
        return self._blocks.numel()

    def kv_blocks(self) -> torch.Tensor:
        return self._blocks[: self._n_blocks]

    def kv_cache_ids(self, cache_group: int = 0, on_device: bool = False) -> torch.Tensor:
        """The whole block table of the sequence (its first ``cur_allocated_blocks`` entries are valid).  The table
        lives on the host; the batch wrapper copies the valid part into its staging buffer per forward, so there is
        no per-sequence device copy to return."""
        if on_device:
            raise ValueError("block tables are staged per batch (RaggedBatchWrapper), not kept per sequence on device")
        return self._blocks

    @property
    def kv_blocks_ptr(self) -> int:
        return self._blocks.data_ptr()

    @property
    def is_offloaded(self) -> bool:
        return self._host_kv is not None

    def detach_kv(self, host_kv: torch.Tensor) -> None:
        """The sequence's KV blocks now live in ``host_kv`` (see ``BlockedKVCache.offload``); it owns no device block."""
        self._host_kv, self._n_blocks = host_kv, 0

    def attach_kv(self, new_ids: torch.Tensor) -> torch.Tensor:
        host, self._host_kv = self._host_kv, None
        self._n_blocks = 0
        self.extend_kv_cache(new_ids)
        return host

    def free_kv_cache(self, free_ids, cache_group: int = 0) -> None:
        raise NotImplementedError("Partial KV-cache freeing is not supported: dense causal attention reads every "
                                  "block of the sequence; use DSStateManager.offload_sequence to release memory")

    def all_block_ids(self) -> torch.Tensor:
        return self.kv_blocks()

    def extend_kv_cache(self, new_ids: torch.Tensor) -> None:
        n = new_ids.numel()
        if self._n_blocks + n > self._blocks.numel():
            raise ValueError("sequence exceeds the maximum number of KV blocks")
        self._blocks[self._n_blocks: self._n_blocks + n] = new_ids
        self._n_blocks += n

    def pre_forward(self, num_tokens: int) -> None:
        self._in_flight_tokens = num_tokens

    def post_forward(self) -> None:
        self._seen_tokens += self._in_flight_tokens
        self._in_flight_tokens = 0
