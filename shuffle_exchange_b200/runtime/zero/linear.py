"""Memory-lean linear for ZeRO-3 (reference: ``runtime/zero/linear.py:50-112``).

``torch.nn.functional.linear`` saves the (gathered, full-size) weight for backward; under ZeRO-3 that pins the
gathered copy of every layer until its backward.  This autograd function saves the weight PARAMETER object
instead (a handle: by the time backward runs, the ZeRO hooks have re-gathered it) together with the input, and
computes ``dX = dY W``, ``dW = dY^T X``, ``db = sum dY`` itself.  On CUDA/bf16 the three products are the tcgen05
GEMM (``ops.functional.gemm``); elsewhere plain matmul.
"""
from __future__ import annotations

import math

import torch
from torch import Tensor
from torch.nn import init
from torch.nn.modules.module import Module
from torch.nn.parameter import Parameter

from ...ops import functional as F


def _mm(a, b, layout):
    if a.is_cuda and a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16:
        return F.gemm(a.contiguous(), b.contiguous(), layout)
    if layout == 0:
        return a @ b.t()
    if layout == 1:
        return a @ b
    return a.t() @ b


class LinearFunctionForZeroStage3(torch.autograd.Function):
    @staticmethod
    def forward(ctx, input, weight, bias=None):
        ctx.save_for_backward(input, weight, bias)
        x2 = input.reshape(-1, input.shape[-1])
        out = _mm(x2, weight, 0)
        if bias is not None:
            out = out + bias
        return out.reshape(*input.shape[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, grad_output):
        input, weight, bias = ctx.saved_tensors
        g2 = grad_output.reshape(-1, grad_output.shape[-1])
        x2 = input.reshape(-1, input.shape[-1])
        grad_input = grad_weight = grad_bias = None
        if ctx.needs_input_grad[0]:
            grad_input = _mm(g2, weight, 1).reshape(input.shape)
        if ctx.needs_input_grad[1]:
            grad_weight = _mm(g2, x2, 2)
        if bias is not None and ctx.needs_input_grad[2]:
            grad_bias = g2.sum(0)
        return grad_input, grad_weight, grad_bias


def zero3_linear_wrap(input, weight, bias=None):
    return LinearFunctionForZeroStage3.apply(input, weight, bias)


class LinearModuleForZeroStage3(Module):
    """Drop-in ``nn.Linear`` whose forward uses :class:`LinearFunctionForZeroStage3`."""
    __constants__ = ["in_features", "out_features"]

    def __init__(self, in_features: int, out_features: int, bias: bool = True, device=None, dtype=None) -> None:
        super().__init__()
        self.in_features, self.out_features = in_features, out_features
        self.weight = Parameter(torch.empty(out_features, in_features, device=device, dtype=dtype))
        if bias:
            self.bias = Parameter(torch.empty(out_features, device=device, dtype=dtype))
        else:
            self.register_parameter("bias", None)
        self.reset_parameters()

    def reset_parameters(self) -> None:
        if self.weight.device.type == "meta":
            return
        init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        if self.bias is not None:
            fan_in, _ = init._calculate_fan_in_and_fan_out(self.weight)
            bound = 1 / math.sqrt(fan_in) if fan_in > 0 else 0
            init.uniform_(self.bias, -bound, bound)

    def forward(self, input: Tensor) -> Tensor:
        return LinearFunctionForZeroStage3.apply(input, self.weight, self.bias)

    def extra_repr(self) -> str:
        return "in_features={}, out_features={}, bias={}".format(self.in_features, self.out_features, self.bias is not None)
