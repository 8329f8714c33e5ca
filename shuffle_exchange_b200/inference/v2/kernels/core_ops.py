"""``core_ops`` of the v2 kernel set (reference ``inference/v2/kernels/core_ops/*``: bias_activations, cuda_layer_norm,
cuda_rms_norm, gated_activations, blas_kernels, cuda_linear) as callables over the framework's sm_100a kernels."""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from ....ops import functional as F
from ..inference_utils import ActivationType

_ACT = {ActivationType.IDENTITY: "identity", ActivationType.RELU: "relu", ActivationType.GELU: "gelu",
        ActivationType.SILU: "silu", ActivationType.GEGLU: "gelu", ActivationType.ReGLU: "relu",
        ActivationType.SiGLU: "silu"}


class CUDABiasActivation:
    """In-place-style ``act(x + bias)`` (reference ``bias_activations/bias_activation.py``)."""
    supported_act_fns = (ActivationType.IDENTITY, ActivationType.GELU, ActivationType.RELU, ActivationType.SILU)

    def __init__(self, channels: int, dtype, act_fn: ActivationType) -> None:
        if ActivationType(act_fn) not in self.supported_act_fns:
            raise ValueError(f"Unsupported activation function: {act_fn}")
        self.act = _ACT[ActivationType(act_fn)]

    def __call__(self, activation: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        if self.act == "identity":
            return activation if bias is None else activation.add_(bias)
        return F.bias_act(activation, bias, self.act)


class CUDAGatedActivation:
    """``act(gate + b) * (up + b)`` halving the last dim; input is ``[.., gate | up]`` (reference
    ``gated_activations/gated_activation.py``)."""
    supported_act_fns = (ActivationType.GEGLU, ActivationType.ReGLU, ActivationType.SiGLU)

    def __init__(self, channels: int, fp_dtype, act_fn: ActivationType) -> None:
        if ActivationType(act_fn) not in self.supported_act_fns:
            raise ValueError(f"Unsupported activation function: {act_fn}")
        self.act = _ACT[ActivationType(act_fn)]

    def __call__(self, output: Optional[torch.Tensor], input_tensor: torch.Tensor,
                 bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        x = input_tensor if bias is None else input_tensor + bias
        y = F.gated_act(x, self.act)
        if output is not None:
            output.copy_(y)
            return output
        return y


class CUDAFPLN:
    def __init__(self, channels: int, fp_dtype, epsilon: float = 1e-5):
        self.eps = epsilon

    def __call__(self, output_z, input_x, gamma, beta):
        y = F.layer_norm(input_x, gamma, beta, self.eps)[0]
        return y if output_z is None else output_z.copy_(y)


class CUDAFPPostLN(CUDAFPLN):
    """``LN(x + residual)``."""

    def __call__(self, output_z, input_x, input_y, gamma, beta):
        y = F.layer_norm(input_x, gamma, beta, self.eps, residual=input_y)[0]
        return y if output_z is None else output_z.copy_(y)


class CUDAFPPreLN(CUDAFPLN):
    """Returns ``(x + residual, LN(x + residual))``."""

    def __call__(self, z_res, z_hid, x_res, x_hid, gamma, beta) -> Tuple[torch.Tensor, torch.Tensor]:
        y, res = F.layer_norm(x_hid, gamma, beta, self.eps, residual=x_res)
        if z_res is not None:
            z_res.copy_(res)
            z_hid.copy_(y)
            return z_res, z_hid
        return res, y


class CUDARMSNorm:
    def __init__(self, channels: int, fp_dtype, epsilon: float = 1e-5):
        self.eps = epsilon

    def __call__(self, output_z, input_x, gamma):
        y = F.rms_norm(input_x, gamma, self.eps)[0]
        return y if output_z is None else output_z.copy_(y)


class CUDARMSPreNorm(CUDARMSNorm):
    def __call__(self, z_res, z_hid, x_res, x_hid, gamma):
        y, res = F.rms_norm(x_hid, gamma, self.eps, residual=x_res)
        if z_res is not None:
            z_res.copy_(res)
            z_hid.copy_(y)
            return z_res, z_hid
        return res, y


class BlasLibLinear:
    """Dense ``y = x W^T`` (reference ``blas_kernels/blas_linear.py`` calls cuBLAS; here the tcgen05 GEMM)."""

    def __init__(self, fp_dtype):
        self.dtype = fp_dtype

    def __call__(self, output: Optional[torch.Tensor], hidden_states: torch.Tensor, weights: torch.Tensor):
        y = F.linear(hidden_states, weights)
        return y if output is None else output.copy_(y)


class CUDAWf6Af16Linear:
    """W6A16 linear (reference ``cuda_linear/cuda_linear.py``: FP6-LLM kernel with 2+4-bit weight fragments).
    Weights arrive FP6-quantised by ``ops.fp_quantizer`` (6 bits + group scales).  Decode-sized inputs run the fused
    weight-only GEMV (codes decoded in registers); larger inputs expand the weight to bf16 and use the tcgen05 GEMM."""

    def __init__(self, group_size: int = 512):
        from ....ops.fp_quantizer import FP_Quantize
        self._q = FP_Quantize(group_size=group_size)

    def quantize(self, weight: torch.Tensor) -> torch.Tensor:
        return self._q.quantize(weight.to(torch.bfloat16).contiguous(), q_bits=6, q_mantisa_bits=2)

    def __call__(self, output, hidden_states, weights_q, shape):
        from ....ops.fp_quantizer import fp_gemv, fp_gemv_ok
        out_f, in_f = int(shape[0]), int(shape[1])
        if fp_gemv_ok(hidden_states, out_f, in_f, self._q.group_size, 6):
            # decode: the 6-bit codes are read once and decoded in registers (weight-bandwidth bound)
            y = fp_gemv(hidden_states, weights_q, out_f, in_f, q_bits=6, group_size=self._q.group_size)
        else:
            # prefill: compute bound — expand to bf16 scratch, tcgen05 GEMM
            w = self._q.dequantize(weights_q, q_bits=6, q_mantisa_bits=2).view(shape).to(hidden_states.dtype)
            y = F.linear(hidden_states, w)
        return y if output is None else output.copy_(y)
