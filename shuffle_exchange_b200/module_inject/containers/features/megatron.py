"""Megatron-LM row order of the fused QKV (reference: ``features/megatron.py``): checkpoints from version 2 on store
``[heads, (q|k|v), d]``; the fused layer wants q|k|v blocks."""
from __future__ import annotations

from ..base import TransformerPolicy


class MegatronContainer:

    def _align_qkv_transposed(self, x, heads=None):
        heads = heads or self.num_attention_heads
        w, _ = TransformerPolicy._deinterleave_qkv(x if x.dim() == 2 else x.view(-1, 1), None, heads)
        return w if x.dim() == 2 else w.view(-1)

    def transpose(self):
        if getattr(self.policy, "megatron_v2", False) and self.attn is not None and self.attn.qkvw is not None \
                and not getattr(self, "_qkv_aligned", False):
            self.attn.qkvw = self._align_qkv_transposed(self.attn.qkvw)
            if self.attn.qkvb is not None:
                self.attn.qkvb = self._align_qkv_transposed(self.attn.qkvb)
            self._qkv_aligned = True
