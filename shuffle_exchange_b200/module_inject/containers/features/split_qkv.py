"""Split q/k/v projections in hybrid mode (reference: ``features/split_qkv.py``): the fused layer's single QKV tensor
is re-assembled from the client's three projections on every refresh."""
from __future__ import annotations

from .hybrid_engine import HybridEngineContainer


class HybridSplitQKVContainer(HybridEngineContainer):

    def set_q_k_v(self):
        """Return the client's (qw, qb, kw, kb, vw, vb) as currently trained."""
        a = self.policy.attention()
        if a.qw is None:
            raise TypeError(f"{type(self.policy).__name__} exposes a fused QKV projection, not split q/k/v")
        self.qw, self.qb, self.kw, self.kb, self.vw, self.vb = a.qw, a.qb, a.kw, a.kb, a.vw, a.vb
        return self.qw, self.qb, self.kw, self.kb, self.vw, self.vb

    def attention_qkv_mp(self, mp_replace=None, reversed_dim=False):
        self.refresh()
