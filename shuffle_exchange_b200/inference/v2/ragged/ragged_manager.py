"""Sequence / KV-cache state manager (reference: ``inference/v2/ragged/ragged_manager.py:19-206``)."""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, Optional

import torch

from .kv_cache import BlockedKVCache, KVCacheConfig
from .sequence_descriptor import DSSequenceDescriptor


@dataclass
class DSStateManagerConfig:
    max_tracked_sequences: int = 2048
    max_ragged_batch_size: int = 768
    max_ragged_sequence_count: int = 512
    max_context: int = 8192
    memory_config_fraction: float = 0.8
    num_kv_blocks: Optional[int] = None
    offload: bool = False
    memory_config: Optional["MemoryConfig"] = None      # reference spelling; overrides the two fields above when set

    def __post_init__(self):
        if isinstance(self.memory_config, dict):
            from .manager_configs import MemoryConfig
            self.memory_config = MemoryConfig(**self.memory_config)
        if self.max_ragged_sequence_count > self.max_tracked_sequences:
            raise ValueError("max_ragged_sequence_count must not exceed max_tracked_sequences")
        if self.max_ragged_sequence_count > self.max_ragged_batch_size:
            raise ValueError("max_ragged_sequence_count must not exceed max_ragged_batch_size")


class DSStateManager:
    def __init__(self, config: DSStateManagerConfig, kv_config: KVCacheConfig, device=None, mp_group=None) -> None:
        self._config, self._kv_config = config, kv_config
        self._seqs: Dict[int, DSSequenceDescriptor] = {}
        num_blocks, reserve = config.num_kv_blocks, None
        mc = config.memory_config
        if mc is not None:
            from .manager_configs import AllocationMode
            if mc.mode == AllocationMode.ALLOCATE:
                num_blocks = int(mc.size)
            else:
                reserve = int(mc.size)
        self._kv_cache = BlockedKVCache(kv_config, num_blocks=num_blocks, memory_fraction=config.memory_config_fraction,
                                        device=device, reserve_bytes=reserve, mp_group=mp_group)
        self.max_blocks_per_seq = (config.max_context + kv_config.block_size - 1) // kv_config.block_size

    def get_sequence(self, uid: int) -> Optional[DSSequenceDescriptor]:
        return self._seqs.get(uid)

    def get_or_create_sequence(self, uid: int) -> DSSequenceDescriptor:
        seq = self._seqs.get(uid)
        if seq is None:
            if len(self._seqs) >= self._config.max_tracked_sequences:
                raise RuntimeError(f"Unable to create tracking slot for sequence {uid}: too many tracked sequences")
            seq = self._seqs[uid] = DSSequenceDescriptor(uid, self.max_blocks_per_seq)
        return seq

    def flush_sequence(self, uid: int) -> None:
        seq = self._seqs.pop(uid, None)
        if seq is None:
            return
        if not seq.is_offloaded:
            self._kv_cache.free(seq.kv_blocks())

    # ------------------------------------------------------------------ suspending sequences to host
    def offload_sequence(self, uid: int) -> int:
        """Move the KV blocks of a tracked, idle sequence to host memory; returns the number of blocks released.  The
        sequence stays tracked (its token count is kept) and must be restored before its next forward."""
        seq = self._seqs[uid]
        if seq.is_offloaded or seq.cur_allocated_blocks == 0:
            return 0
        if seq.in_flight_tokens:
            raise RuntimeError(f"sequence {uid} has tokens in flight")
        n = seq.cur_allocated_blocks
        seq.detach_kv(self._kv_cache.offload(seq.kv_blocks()))
        return n

    def restore_sequence(self, uid: int) -> int:
        """Bring an offloaded sequence's KV blocks back (into whichever blocks are free now)."""
        seq = self._seqs[uid]
        if not seq.is_offloaded:
            return 0
        n = seq._host_kv.shape[1]
        if n > self.free_blocks:
            raise ValueError(f"restoring sequence {uid} needs {n} KV blocks, {self.free_blocks} are free")
        seq.attach_kv(self._kv_cache.restore(seq._host_kv))
        return n

    @property
    def n_kv_cache_groups(self) -> int:
        return 1

    def allocate_blocks(self, n_blocks: int) -> torch.Tensor:
        return self._kv_cache.reserve(n_blocks)

    def get_cache(self, layer: int) -> torch.Tensor:
        return self._kv_cache.get_cache(layer)

    @property
    def free_blocks(self) -> int:
        return self._kv_cache.free_blocks

    @property
    def kv_block_size(self) -> int:
        return self._kv_config.block_size

    @property
    def n_tracked_sequences(self) -> int:
        return len(self._seqs)

    @property
    def tracked_sequences(self):
        return self._seqs
