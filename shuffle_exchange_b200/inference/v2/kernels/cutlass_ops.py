"""``cutlass_ops`` of the v2 kernel set (reference ``inference/v2/kernels/cutlass_ops/{mixed_gemm,moe_gemm}``: the
reference forwards to the external ``deepspeed-kernels`` CUTLASS runners — grouped GEMM over experts and W4/W8-A16
GEMM).  Here: per-expert tcgen05 GEMMs over the contiguous row ranges produced by ``moe_scatter`` (the cumulative
``total_rows_before_expert`` vector is the same interface), and weight-only INT4/INT8 GEMM as dequantise + GEMM."""
from __future__ import annotations

from typing import Optional

import torch

from ....ops import functional as F
from ....ops.quantizer import dequantize
from ..inference_utils import ActivationType
from .core_ops import _ACT


def _epilogue(y: torch.Tensor, bias: Optional[torch.Tensor], act: ActivationType) -> torch.Tensor:
    act = ActivationType(act)
    if act == ActivationType.IDENTITY:
        return y if bias is None else y + bias
    if act in (ActivationType.GEGLU, ActivationType.ReGLU, ActivationType.SiGLU):
        return F.gated_act(y if bias is None else y + bias, _ACT[act])
    return F.bias_act(y, bias, _ACT[act])


class MoEGEMM:
    """``ordered_output[rows of expert e] = act(ordered_input[rows of e] @ W[e]^T + b[e])`` for every expert;
    ``total_rows_before_expert`` is the inclusive cumulative row count (int64/int32 ``[E]``)."""
    supported_dtypes = (torch.float16, torch.bfloat16, torch.float32)
    supported_act_fns = (ActivationType.GELU, ActivationType.SILU, ActivationType.RELU, ActivationType.IDENTITY,
                         ActivationType.GEGLU, ActivationType.SiGLU, ActivationType.ReGLU)

    def __init__(self, fp_dtype, act_fn: ActivationType) -> None:
        if fp_dtype not in self.supported_dtypes:
            raise ValueError(f"Unsupported data type: {fp_dtype}")
        if ActivationType(act_fn) not in self.supported_act_fns:
            raise ValueError(f"Unsupported activation function: {act_fn}")
        self.act_fn = ActivationType(act_fn)

    def __call__(self, ordered_output: Optional[torch.Tensor], ordered_input: torch.Tensor, weights: torch.Tensor,
                 total_rows_before_expert: torch.Tensor, biases: Optional[torch.Tensor] = None) -> torch.Tensor:
        ends = total_rows_before_expert.tolist()
        pieces, lo = [], 0
        for e, hi in enumerate(ends):
            if hi > lo:
                y = _epilogue(F.linear(ordered_input[lo:hi], weights[e]), None if biases is None else biases[e],
                              self.act_fn)
                if ordered_output is not None:
                    ordered_output[lo:hi] = y
                else:
                    pieces.append(y)
            lo = hi
        if ordered_output is not None:
            return ordered_output
        return torch.cat(pieces) if pieces else ordered_input.new_empty(0, weights.shape[1])


class MixedGEMM:
    """Weight-only quantised GEMM: ``act(x @ dequant(Wq)^T + b)`` with INT4/INT8 group-quantised weights."""
    supported_act_fns = MoEGEMM.supported_act_fns

    def __init__(self, fp_dtype, act_fn: ActivationType, num_bits: int, groups: int = 64, symmetric: bool = True):
        if num_bits not in (4, 8):
            raise ValueError(f"Unsupported num_bits: {num_bits}")
        self.fp_dtype, self.act_fn = fp_dtype, ActivationType(act_fn)
        self.num_bits, self.groups, self.symmetric = num_bits, groups, symmetric

    def __call__(self, output: Optional[torch.Tensor], hidden_states: torch.Tensor, weights: torch.Tensor,
                 scales: torch.Tensor, shape, biases: Optional[torch.Tensor] = None) -> torch.Tensor:
        w = dequantize(weights, scales, self.groups, self.num_bits, self.symmetric, hidden_states.dtype).view(shape)
        y = _epilogue(F.linear(hidden_states, w), biases, self.act_fn)
        return y if output is None else output.copy_(y)


class MixedMoEGEMM(MoEGEMM):
    """Grouped GEMM over experts with INT4/INT8 weights (reference ``moe_gemm/mixed_moe_gemm.py``)."""

    def __init__(self, fp_dtype, act_fn: ActivationType, num_bits: int, groups: int = 64, symmetric: bool = True):
        super().__init__(fp_dtype, act_fn)
        if num_bits not in (4, 8):
            raise ValueError(f"Unsupported num_bits: {num_bits}")
        self.num_bits, self.groups, self.symmetric = num_bits, groups, symmetric

    def __call__(self, ordered_output, ordered_input, weights, scales, shape, total_rows_before_expert, biases=None):
        dense = torch.stack([dequantize(weights[e], scales[e], self.groups, self.num_bits, self.symmetric,
                                        ordered_input.dtype).view(shape) for e in range(weights.shape[0])])
        return super().__call__(ordered_output, ordered_input, dense, total_rows_before_expert, biases)
