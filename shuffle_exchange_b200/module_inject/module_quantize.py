"""Post-training INT8 quantisation of transformer layer weights in place (reference: ``module_inject/module_quantize.py``).

``quantize_transformer_layer(BertLayer, model, megatron=False, preln=False)`` converts the 2-D projection weights of
every ``orig_layer_impl`` instance to int8 values stored in the layer's original dtype container (symmetric, one scale
per output-feature group, as the reference's ``quantize_weight`` does with ``torch.round(x / scale)``)."""
from __future__ import annotations

import torch
import torch.nn as nn


def _quantize_weight(weight: torch.Tensor, groups: int = 1, bits: int = 8):
    w = weight.data.float()
    flat = w.reshape(groups, -1)
    scale = flat.abs().amax(dim=1, keepdim=True).clamp_min(1e-8) / float((1 << (bits - 1)) - 1)
    q = torch.round(flat / scale).clamp_(-(1 << (bits - 1)), (1 << (bits - 1)) - 1)
    return q.reshape(w.shape).to(torch.int8), scale.reshape(-1)


def _quantize_module(model: nn.Module, policies: dict):
    for name, child in list(model.named_children()):
        fn = policies.get(type(child))
        if fn is not None:
            setattr(model, name, fn(child))
        else:
            _quantize_module(child, policies)
    return model


def quantize_module(model, orig_class, quantize_fn):
    return _quantize_module(model, {orig_class: quantize_fn})


def quantize_transformer_layer(orig_layer_impl, model, megatron: bool = False, preln: bool = False, groups: int = 1):
    """Quantise attention and MLP projection weights of every ``orig_layer_impl``; each converted ``nn.Linear`` keeps
    ``weight`` as int8 (``requires_grad=False``) and gains ``weight_scale`` so ``dequantized_weight()`` is exact."""

    def quantize_fn(layer):
        for lin in (m for m in layer.modules() if isinstance(m, nn.Linear)):
            q, scale = _quantize_weight(lin.weight, groups)
            lin.weight = nn.Parameter(q, requires_grad=False)
            lin.register_buffer("weight_scale", scale.to(lin.bias.dtype if lin.bias is not None else torch.float32))
            lin.dequantized_weight = (lambda l=lin: (l.weight.float().reshape(groups, -1) *
                                                     l.weight_scale.float()[:, None]).reshape(l.weight.shape))
            lin.forward = (lambda x, l=lin: torch.nn.functional.linear(x, l.dequantized_weight().to(x.dtype), l.bias))
        return layer

    return quantize_module(model, orig_layer_impl, quantize_fn)
