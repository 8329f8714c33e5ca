"""CUDA-graphed UNet wrapper (reference ``model_implementations/diffusers/unet.py``): channels-last weights, one
captured graph for the denoising step, ``forward(sample, timestamp, encoder_hidden_states, return_dict=True, ...)``."""
from __future__ import annotations

import torch

from ..features.graphed_module import GraphedModule


class DSUNet(GraphedModule):

    def __init__(self, unet, enable_cuda_graph=True):
        super().__init__(enable_cuda_graph=enable_cuda_graph)
        self.unet = unet
        self.in_channels = getattr(unet, "in_channels", None)
        self.device = getattr(unet, "device", None)
        self.dtype = getattr(unet, "dtype", None)
        self.config = getattr(unet, "config", None)
        self.fwd_count = 0
        self.unet.requires_grad_(requires_grad=False)
        self.unet.to(memory_format=torch.channels_last)
        self.cuda_graph_created = False

    def forward(self, sample, timestamp, encoder_hidden_states, return_dict=True, cross_attention_kwargs=None,
                timestep_cond=None, added_cond_kwargs=None):
        if cross_attention_kwargs:
            raise NotImplementedError("cross_attention_kwargs is not supported by the graphed UNet")
        kw = {}
        if timestep_cond is not None:
            kw["timestep_cond"] = timestep_cond
        if added_cond_kwargs is not None:
            kw["added_cond_kwargs"] = added_cond_kwargs
        if self.enable_cuda_graph and kw:
            # dict-valued conditioning changes the captured signature: run eagerly
            return self._forward(sample, timestamp, encoder_hidden_states, return_dict, **kw)
        out = self._run("forward", self._forward, sample, timestamp, encoder_hidden_states, return_dict, **kw)
        self.cuda_graph_created = "forward" in self._graphs
        return out

    def _forward(self, sample, timestamp, encoder_hidden_states, return_dict=True, **kw):
        return self.unet(sample, timestamp, encoder_hidden_states, return_dict=return_dict, **kw)
