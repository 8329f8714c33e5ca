"""Quantised drop-in layers (reference ``deepspeed/inference/quantization/layers.py:20-114``).

The weight of the wrapped layer is replaced by one packed byte parameter (codes | scales | mins); ``forward``
dequantises into the activation dtype and runs the normal op — on B200 the dequantised operand feeds the tcgen05 GEMM,
so the win is HBM capacity/bandwidth for the resident weights, which is what post-init quantisation is for."""
from __future__ import annotations

from typing import Callable, Dict, Tuple

import torch
import torch.nn.functional as F
from torch import Tensor, nn

from .utils import DeQuantizer, Quantizer, concat_to_compat_param

quantized_weight_registry: Dict[int, nn.Parameter] = {}
is_zero3_enabled = False


def get_quantize_weight_fn(quantizer: Quantizer, pre_quant_weight: nn.Parameter) -> Callable:
    def func() -> Tuple[nn.Parameter, Callable]:
        codes, scale, mn = quantizer.quantize(pre_quant_weight.data)
        return concat_to_compat_param(codes, scale, mn)
    return func


def get_quantized_weight_wrapper(model, pre_quant_weight: nn.Parameter, quantize_weight_fn: Callable):
    """Tied weights (e.g. embedding <-> lm head) are quantised once and shared."""
    key = pre_quant_weight.data_ptr() if pre_quant_weight.numel() else id(pre_quant_weight)
    if key in quantized_weight_registry:
        packed, deconcat = quantized_weight_registry[key]
    else:
        packed, deconcat = quantize_weight_fn()
        quantized_weight_registry[key] = (packed, deconcat)
    model._deconcat = deconcat
    return packed


class QuantizedLinear(nn.Linear):

    def __init__(self, config: Dict, pre_quant_layer: nn.Linear) -> None:
        super().__init__(in_features=pre_quant_layer.in_features, out_features=pre_quant_layer.out_features,
                         bias=pre_quant_layer.bias is not None, device="meta", dtype=pre_quant_layer.weight.dtype)
        self.config = config
        self.quantizer = Quantizer(config=config)
        self.bias = pre_quant_layer.bias
        self.compute_dtype = pre_quant_layer.weight.dtype
        self.weight = get_quantized_weight_wrapper(self, pre_quant_layer.weight,
                                                   get_quantize_weight_fn(self.quantizer, pre_quant_layer.weight))
        self.weight.dequantizer = DeQuantizer(config, self.compute_dtype)

    def dequantized_weight(self) -> Tensor:
        codes, scale, mn = self._deconcat(self.weight.data)
        return self.weight.dequantizer.dequantize(codes, scale, mn)

    def forward(self, input: Tensor) -> Tensor:
        return F.linear(input, self.dequantized_weight().to(input.dtype), self.bias)


class QuantizedEmbedding(nn.Embedding):

    def __init__(self, config: Dict, pre_quant_layer: nn.Embedding) -> None:
        super().__init__(num_embeddings=pre_quant_layer.num_embeddings, embedding_dim=pre_quant_layer.embedding_dim,
                         padding_idx=pre_quant_layer.padding_idx, max_norm=pre_quant_layer.max_norm,
                         norm_type=pre_quant_layer.norm_type, scale_grad_by_freq=pre_quant_layer.scale_grad_by_freq,
                         sparse=pre_quant_layer.sparse, _weight=pre_quant_layer.weight, device=None,
                         dtype=pre_quant_layer.weight.dtype)
        assert pre_quant_layer.max_norm is None, "max_norm is not supported with a quantised table"
        self.config = config
        self.quantizer = Quantizer(config=config)
        self.compute_dtype = pre_quant_layer.weight.dtype
        self.weight = get_quantized_weight_wrapper(self, pre_quant_layer.weight,
                                                   get_quantize_weight_fn(self.quantizer, pre_quant_layer.weight))
        self.weight.dequantizer = DeQuantizer(config, self.compute_dtype)

    def dequantized_weight(self) -> Tensor:
        codes, scale, mn = self._deconcat(self.weight.data)
        return self.weight.dequantizer.dequantize(codes, scale, mn)

    def forward(self, input: Tensor) -> Tensor:
        return F.embedding(input, self.dequantized_weight(), self.padding_idx, self.max_norm, self.norm_type,
                           self.scale_grad_by_freq, self.sparse)


QUANTIZATION_LAYER_MAPPINGS = {nn.Linear: QuantizedLinear, nn.Embedding: QuantizedEmbedding}
