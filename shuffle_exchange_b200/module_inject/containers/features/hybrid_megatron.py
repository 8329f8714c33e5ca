"""Megatron fused QKV in hybrid mode (reference: ``features/hybrid_megatron.py``)."""
from __future__ import annotations

from .hybrid_engine import HybridEngineContainer
from .megatron import MegatronContainer


class HybridMegatronContainer(MegatronContainer, HybridEngineContainer):

    def transform_for_inference(self):
        self._qkv_aligned = False
        self.refresh()
