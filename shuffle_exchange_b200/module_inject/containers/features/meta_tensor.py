"""Models created on the ``meta`` device (no storage) get their layer weights straight from checkpoint state dicts at
injection time (reference: ``features/meta_tensor.py``)."""
from __future__ import annotations

import torch


class MetaTensorContainer:
    """Mix-in for :class:`BaseTransformerContainer`.  ``load_params(sd, prefix)`` materialises the client layer from
    ``sd`` (keys ``prefix + <layer-relative name>``), after which the ordinary ``initialize_tensors`` path applies."""

    @property
    def is_meta(self) -> bool:
        return any(p.device.type == "meta" for p in self.child.parameters())

    def load_params(self, sd, prefix: str = "", mp_replace=None, weight_quantizer=None):
        from ...load_checkpoint import load_model_with_checkpoint
        missing = load_model_with_checkpoint(self.child, sd, mp_replace=None, weight_quantizer=weight_quantizer,
                                             prefixes=(prefix,))
        still_meta = [n for n, p in self.child.named_parameters() if p.device.type == "meta"]
        if still_meta:
            raise KeyError(f"checkpoint lacks {still_meta[:4]}{' ...' if len(still_meta) > 4 else ''} under prefix {prefix!r}")
        return missing

    def initialize_tensors(self, enable_training: bool = False):
        if self.is_meta:
            raise RuntimeError("layer is still on the meta device: call load_params(state_dict, prefix) first")
        return super().initialize_tensors()
