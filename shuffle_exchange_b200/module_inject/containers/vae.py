"""diffusers ``AutoencoderKL`` (reference: ``containers/vae.py``): wrapped in :class:`DSVAE` (graphed encode/decode)."""
from __future__ import annotations

from .base import TransformerPolicy


class VAEPolicy(TransformerPolicy):
    _orig_layer_class_name = "AutoencoderKL"

    @classmethod
    def match(cls, module):
        return type(module).__name__ == "AutoencoderKL"

    def apply(self, module, enable_cuda_graph=True):
        from ...model_implementations.diffusers.vae import DSVAE
        return DSVAE(module, enable_cuda_graph=enable_cuda_graph)
