"""diffusers ``UNet2DConditionModel`` (reference: ``containers/unet.py``): the whole model is wrapped in a CUDA-graph
replaying :class:`DSUNet`; its attention blocks are the generic-injection targets."""
from __future__ import annotations

from .base import TransformerPolicy


class UNetPolicy(TransformerPolicy):
    _orig_layer_class_name = "UNet2DConditionModel"

    @classmethod
    def match(cls, module):
        return type(module).__name__ == "UNet2DConditionModel"

    def apply(self, module, enable_cuda_graph=True):
        from ...model_implementations.diffusers.unet import DSUNet
        return DSUNet(module, enable_cuda_graph=enable_cuda_graph)

    def attention(self, client_module=None):
        """(qkvw or None, qw, kw, vw, ow, ob, heads) of a diffusers attention block (self-attention fuses q/k/v)."""
        c = client_module if client_module is not None else self.client_module
        qw, kw, vw = c.to_q.weight.data, c.to_k.weight.data, c.to_v.weight.data
        fused = None
        if qw.shape[1] == kw.shape[1]:
            import torch
            fused = torch.cat([qw, kw, vw], dim=0)
        return fused, qw, kw, vw, c.to_out[0].weight.data, c.to_out[0].bias.data, c.heads
