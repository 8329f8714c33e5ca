"""Quantised drop-in layers (reference ``deepspeed/inference/quantization/layers.py:20-114``).

The weight of the wrapped layer is replaced by one packed byte parameter (codes | scales | mins); ``forward``
dequantises into the activation dtype and runs the normal op — on B200 the dequantised operand feeds the tcgen05 GEMM,
so the win is HBM capacity/bandwidth for the resident weights, which is what post-init quantisation is for.  At decode
sizes (at most ``GEMV_MAX_ROWS`` activation rows) a linear layer grouped along its input dimension skips the
dequantised copy: ``csrc/quantize.cu: int_gemv_kernel`` reads the codes once and applies scale / min per group in
registers (the reference dequantises for every batch size)."""
from __future__ import annotations

from typing import Callable, Dict, Tuple

import torch
import torch.nn.functional as F
from torch import Tensor, nn

from ...native import lib as _lib
from .utils import DeQuantizer, Quantizer, concat_to_compat_param

_lib.register({"sxb_int_gemv": "ppppp" "iiiii" "ll" "p"})

quantized_weight_registry: Dict[tuple, Tuple[nn.Parameter, Callable]] = {}
is_zero3_enabled = False
GEMV_MAX_ROWS = 4


def int_gemv_ok(x: Tensor, codes: Tensor, scale: Tensor, config: Dict, in_features: int) -> bool:
    """Whether :func:`int_gemv` takes this problem: bf16 CUDA activations of at most ``GEMV_MAX_ROWS`` rows, groups
    along the input dimension of a 2-D weight, aligned sizes."""
    gs = int(config["group_size"])
    rows = x.numel() // max(1, x.shape[-1])
    return (x.is_cuda and x.dtype == torch.bfloat16 and scale.dtype == torch.bfloat16 and codes.dim() == 2
            and 0 < rows <= GEMV_MAX_ROWS and config["group_dim"] in (1, -1) and x.shape[-1] == in_features
            and in_features % 256 == 0 and gs % 8 == 0 and in_features % gs == 0 and codes.data_ptr() % 8 == 0)


def int_gemv(x: Tensor, codes: Tensor, scale: Tensor, mn: Tensor, out_features: int, in_features: int, num_bits: int,
             group_size: int) -> Tensor:
    """``x @ W^T`` straight from the packed codes (``W = codes / scale + min`` per group, never materialised)."""
    x2 = x.reshape(-1, in_features)
    if not x2.is_contiguous():
        x2 = x2.contiguous()
    y = torch.empty(x2.shape[0], out_features, dtype=torch.bfloat16, device=x.device)
    _lib.call("sxb_int_gemv", x2.data_ptr(), codes.data_ptr(), scale.data_ptr(), mn.data_ptr(), y.data_ptr(),
              x2.shape[0], out_features, in_features, int(num_bits), int(group_size), x2.stride(0), y.stride(0),
              _lib.stream_ptr())
    return y.view(*x.shape[:-1], out_features)


def get_quantize_weight_fn(quantizer: Quantizer, pre_quant_weight: nn.Parameter) -> Callable:
    def func() -> Tuple[nn.Parameter, Callable]:
        codes, scale, mn = quantizer.quantize(pre_quant_weight.data)
        return concat_to_compat_param(codes, scale, mn)
    return func


def get_quantized_weight_wrapper(model, pre_quant_weight: nn.Parameter, quantize_weight_fn: Callable):
    """Tied weights (e.g. embedding <-> lm head) are quantised once and shared."""
    cfg = getattr(model, "config", None) or {}
    # the storage address alone is not an identity (a freed weight's address is reused by the next allocation):
    # shape and quantisation settings are part of the key
    key = (pre_quant_weight.data_ptr() if pre_quant_weight.numel() else id(pre_quant_weight),
           tuple(pre_quant_weight.shape), cfg.get("num_bits"), cfg.get("group_size"), cfg.get("group_dim"))
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
        codes, scale, mn = self._deconcat(self.weight.data)
        if int_gemv_ok(input, codes, scale, self.config, self.in_features):
            y = int_gemv(input, codes, scale, mn, self.out_features, self.in_features, self.config["num_bits"],
                         self.config["group_size"])
            return y if self.bias is None else y + self.bias.to(y.dtype)
        w = self.weight.dequantizer.dequantize(codes, scale, mn)
        return F.linear(input, w.to(input.dtype), self.bias)


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
