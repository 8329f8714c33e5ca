"""CUDA-graphed VAE wrapper (reference ``model_implementations/diffusers/vae.py``): separate graphs for ``decode``,
``encode`` and the full ``forward``."""
from __future__ import annotations

from ..features.graphed_module import GraphedModule


class DSVAE(GraphedModule):

    def __init__(self, vae, enable_cuda_graph=True):
        super().__init__(enable_cuda_graph=enable_cuda_graph)
        self.vae = vae
        self.config = getattr(vae, "config", None)
        self.device = getattr(vae, "device", None)
        self.dtype = getattr(vae, "dtype", None)
        self.vae.requires_grad_(requires_grad=False)
        self.decoder_cuda_graph_created = False
        self.encoder_cuda_graph_created = False
        self.all_cuda_graph_created = False

    def _decode(self, x, return_dict=True, generator=None):
        return self.vae.decode(x, return_dict=return_dict)

    def decode(self, *inputs, **kwargs):
        out = self._run("decode", self._decode, *inputs, **kwargs)
        self.decoder_cuda_graph_created = "decode" in self._graphs
        return out

    def _encode(self, x, return_dict=True):
        return self.vae.encode(x, return_dict=return_dict)

    def encode(self, *inputs, **kwargs):
        out = self._run("encode", self._encode, *inputs, **kwargs)
        self.encoder_cuda_graph_created = "encode" in self._graphs
        return out

    def _forward(self, sample, timestamp=None, encoder_hidden_states=None, return_dict=True):
        return self.vae(sample, timestamp, encoder_hidden_states, return_dict)

    def forward(self, *inputs, **kwargs):
        out = self._run("forward", self._forward, *inputs, **kwargs)
        self.all_cuda_graph_created = "forward" in self._graphs
        return out
