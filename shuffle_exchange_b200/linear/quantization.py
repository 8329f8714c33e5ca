"""Weight-only quantised parameters and linear layer.

Reference: ``deepspeed/linear/quantization.py:18-147`` — ``QuantizedParameter`` quantises itself the moment it lands
on the accelerator (``.to(device)`` / ``.cuda()``), keeps the payload as ``uint8`` and dequantises on demand;
``QuantizedLinear`` dequantises then calls the dense linear.  Quantiser: the FP8/FP6/FP4 group-scaled kernels of
``csrc/quantize.cu`` (``ops.fp_quantizer``).
"""
from __future__ import annotations

import copy
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as TF
from torch.nn.modules.linear import Linear as _TorchLinear   # immune to linear.Init's nn.Linear patch

from ..ops.fp_quantizer import FP_Quantize
from .config import QuantizationConfig


class QuantizedParameter(nn.Parameter):
    """Parameter that stores group-scaled low-precision bytes once it is on a CUDA device."""

    def __new__(cls, data: Optional[torch.Tensor] = None, requires_grad: bool = False,
                quantization_config: QuantizationConfig = None, quantizer=None):
        if requires_grad:
            raise ValueError("requires_grad=True is not supported with QuantizedParameter")
        if data is None:
            data = torch.empty(0)
        self = torch.Tensor._make_subclass(cls, data, requires_grad)
        self.quantization_config = QuantizationConfig() if quantization_config is None else quantization_config
        self.quantizer = quantizer if quantizer is not None else FP_Quantize(group_size=self.quantization_config.group_size)
        self._ensure_quantized(self)
        return self

    def _ensure_quantized(self, tensor: torch.Tensor):
        if tensor.device.type == "cuda" and tensor.dtype != self.quantization_config.q_dtype:
            with torch.no_grad():
                self.quantizer.orig_dtype = None
                payload = self.quantizer.quantize(tensor.data.contiguous(), q_bits=self.quantization_config.q_bits,
                                                  q_mantisa_bits=self.quantization_config.mantissa_bits)
                tensor.data = payload
            assert tensor.dtype == self.quantization_config.q_dtype

    def dequantized(self) -> torch.Tensor:
        if self.device.type == "cuda" and self.dtype == self.quantization_config.q_dtype:
            with torch.no_grad():
                return self.quantizer.dequantize(self.data, q_bits=self.quantization_config.q_bits,
                                                 q_mantisa_bits=self.quantization_config.mantissa_bits)
        return self.data

    def offload(self, revert=False):
        self.data = self.data.to("cuda" if revert else "cpu", non_blocking=True)
        self.quantizer.to("cuda" if revert else "cpu")

    def __getstate__(self):
        state = self.__dict__.copy()
        state["data"] = self.data
        state["quantization_config"] = self.quantization_config
        state["requires_grad"] = self.requires_grad
        return state

    def __setstate__(self, state):
        self.quantizer = state["quantizer"]
        self.quantization_config = state["quantization_config"]
        self.data = state["data"]
        self.requires_grad = state["requires_grad"]

    def __deepcopy__(self, memo):
        new = type(self).__new__(type(self))
        st = self.__getstate__()
        new.__setstate__(st)
        new.quantizer = copy.deepcopy(st["quantizer"])
        new.data = copy.deepcopy(st["data"])
        return new

    def __copy__(self):
        new = type(self).__new__(type(self))
        new.__setstate__(self.__getstate__())
        return new

    def cuda(self, device=None, non_blocking=False):
        return self.to(device="cuda" if device is None else device, non_blocking=non_blocking)

    def to(self, *args, **kwargs):
        moved = super().to(*args, **kwargs)
        out = QuantizedParameter.__new__(QuantizedParameter, moved.data, False, self.quantization_config, self.quantizer)
        return out


class QuantizedLinear(_TorchLinear):
    """``nn.Linear`` whose weight is a :class:`QuantizedParameter` (frozen); forward = dequantise + GEMM."""

    def __init__(self, input_dim: int, output_dim: int, bias: bool = False,
                 quantization_config: QuantizationConfig = None, dtype=torch.bfloat16):
        super().__init__(input_dim, output_dim, bias=bias, dtype=dtype)
        assert dtype == torch.bfloat16, "currently only supports bfloat16 dtype"
        self.weight = QuantizedParameter(self.weight.data, quantization_config=quantization_config)

    def forward(self, input: torch.Tensor) -> torch.Tensor:
        w = self.weight.dequantized().view(self.out_features, self.in_features).to(input.dtype)
        return TF.linear(input, w, self.bias)
