"""Configuration objects of the ragged state manager (reference: ``inference/v2/ragged/manager_configs.py:15-183``).

``KVCacheConfig`` and ``DSStateManagerConfig`` are defined next to the classes they configure (``kv_cache.py``,
``ragged_manager.py``) and re-exported here under the reference's module name; ``MemoryConfig`` adds the reference's two
allocation modes on top of them.
"""
from __future__ import annotations

from dataclasses import dataclass
from enum import Enum

from .kv_cache import KVCacheConfig  # noqa: F401


class KVCacheType(Enum):
    DENSE = "dense"      # every layer attends over the whole sequence
    LOCAL = "local"      # sliding-window layers keep only the last ``window`` tokens

    @staticmethod
    def from_str(s: str) -> "KVCacheType":
        try:
            return KVCacheType(s.lower())
        except ValueError:
            raise ValueError(f"Unknown KV cache type {s!r}; expected one of {[t.value for t in KVCacheType]}")


class AllocationMode(Enum):
    RESERVE = "reserve"      # fill free HBM with KV blocks but leave ``size`` bytes for everything else
    ALLOCATE = "allocate"    # exactly ``size`` blocks


@dataclass
class MemoryConfig:
    mode: AllocationMode = AllocationMode.RESERVE
    size: int = 1_000_000_000

    def __post_init__(self):
        if isinstance(self.mode, str):
            self.mode = AllocationMode(self.mode.lower())
        if int(self.size) <= 0:
            raise ValueError("MemoryConfig.size must be positive")


from .ragged_manager import DSStateManagerConfig  # noqa: E402,F401
