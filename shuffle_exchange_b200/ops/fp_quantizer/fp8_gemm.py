"""Weight-only FP8 GEMM: ``y = x @ dequant(W_q)``.

Reference: ``deepspeed/ops/fp_quantizer/fp8_gemm.py:23-28`` (dequantise + matmul fallback) and the Triton kernel
``fp8_gemm_triton.py:17-165``.  Here the weight is expanded by the grouped dequantise kernel into a bf16 scratch
tile-set that stays L2-resident for moderate sizes and is consumed by the tcgen05 GEMM (``csrc/gemm_tcgen05.cu``);
a fully fused dequant-in-the-producer variant is the planned next step.
"""
import torch

from .. import functional as F
from .quantize import fp_dequantize


def matmul_fp8(inp: torch.Tensor, weight: torch.Tensor, scale: torch.Tensor, quantization_group_size: int,
               quantizer=None, q_bits: int = 8) -> torch.Tensor:
    """``inp`` [..., K] bf16; ``weight`` uint8 payload of a [K, N] matrix quantised in groups along its flattening
    order; ``scale`` uint8 [groups, 4] (fp32 bytes) or None when the scales are already appended to ``weight``."""
    K = inp.shape[-1]
    payload = weight.reshape(-1, quantization_group_size * q_bits // 8) if scale is not None else weight
    if scale is not None:
        q = torch.cat([payload, scale.reshape(-1, 4)], dim=1).contiguous()
    else:
        q = payload
    w = fp_dequantize(q, quantization_group_size, q_bits, inp.dtype).reshape(K, -1)
    x2 = inp.reshape(-1, K)
    if x2.is_cuda:
        y = F.gemm(x2.contiguous(), w, 1)          # NN layout: [M, K] x [K, N]
    else:
        y = x2 @ w
    return y.reshape(*inp.shape[:-1], w.shape[1])
