"""Gated MLPs in hybrid mode (reference: ``features/gated_mlp.py``): up and gate projections stay separate in the
client and are re-fused (up | gate) on refresh."""
from __future__ import annotations

from .hybrid_engine import HybridEngineContainer


class HybridGatedMLPContainer(HybridEngineContainer):

    def set_mlp_gate(self):
        m = self.policy.mlp()
        if m.up_w is None:
            raise TypeError(f"{type(self.policy).__name__} has no gated MLP")
        self.inter_up_w, self.inter_gate_w = m.up_w, m.gate_w
        return self.inter_up_w, self.inter_gate_w

    def mlp_inter_mp(self, mp_replace=None, reversed_dim=False):
        self.refresh()
