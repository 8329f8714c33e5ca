"""Ragged-batch state: paged KV cache, block allocator, sequence descriptors, the batch staging buffer
(reference: ``inference/v2/ragged/__init__.py``)."""
from .blocked_allocator import BlockedAllocator
from .kv_cache import BlockedKVCache, split_kv
from .manager_configs import AllocationMode, DSStateManagerConfig, KVCacheConfig, KVCacheType, MemoryConfig
from .ragged_manager import DSStateManager
from .ragged_wrapper import RaggedBatchWrapper
from .sequence_descriptor import BaseSequenceDescriptor, DSSequenceDescriptor, PlaceholderSequenceDescriptor
