"""Evoformer (AlphaFold-style) attention with a mask bias and a pair bias.

Reference: ``deepspeed/ops/deepspeed4science/evoformer_attn.py:15-106`` — ``DS4Sci_EvoformerAttention(Q, K, V, biases)``
with ``Q/K/V [B, N, L, H, D]``, ``bias1 [B, N, 1, 1, L]`` (sequence mask) and ``bias2 [B, 1, H, L, L]`` (pair bias), both
optional, gradients for both biases.  The reference calls a CUTLASS kernel that is not in its tree.

B200 path: without biases the rows go straight through the tcgen05 flash-attention kernel (``[B*N]`` as batch); a
sequence mask alone (``bias1`` that needs no gradient) rides along as the kernels' additive per-key bias; with a pair
bias the ``L x L`` scores of the (short, L <= a few thousand) MSA rows are produced by one batched tensor-core GEMM and
the bias add + scale + softmax (forward) / softmax gradient + bias gradients (backward) run in a single fused kernel each
(``csrc/softmax_dropout.cu``), both biases read through broadcast strides — never expanded in memory.
"""
from __future__ import annotations

import math
from typing import List, Optional

import torch

from .. import functional as F
from ..fused_softmax import biased_softmax


def _evoformer_core(q, k, v, bias1, bias2):
    """q/k/v: [B, N, L, H, D]; returns [B, N, L, H, D]."""
    B, N, L, H, D = q.shape
    scale = 1.0 / math.sqrt(D)
    if bias1 is None and bias2 is None and q.is_cuda and q.dtype == torch.bfloat16 and D in (64, 128):
        out = F.flash_attention(q.reshape(B * N, L, H, D), k.reshape(B * N, L, H, D), v.reshape(B * N, L, H, D),
                                causal=False, scale=scale)
        return out.view(B, N, L, H, D)
    # mask only (no pair bias) and no gradient wanted for the mask: it is a per-key bias of the [B*N] rows, which the
    # flash kernels add to the scores themselves (forward and backward)
    if bias2 is None and bias1 is not None and not bias1.requires_grad \
            and F.key_bias_supported(q.reshape(B * N, L, H, D), bias1.reshape(B * N, L)):
        out = F.flash_attention_key_bias(q.reshape(B * N, L, H, D), k.reshape(B * N, L, H, D),
                                         v.reshape(B * N, L, H, D), bias1.reshape(B * N, L), causal=False, scale=scale)
        return out.view(B, N, L, H, D)
    outs = []
    for b in range(B):   # the pair bias is shared by the N rows of one batch element: one fused call per element
        qb = q[b].permute(0, 2, 1, 3)                      # [N, H, L, D]
        kb = k[b].permute(0, 2, 1, 3)
        vb = v[b].permute(0, 2, 1, 3)
        scores = torch.matmul(qb, kb.transpose(-1, -2))    # [N, H, L, L]
        b1 = bias1[b] if bias1 is not None else None       # [N, 1, 1, L]
        b2 = bias2[b] if bias2 is not None else None       # [1, H, L, L]
        probs = biased_softmax(scores, b1, b2, scale)
        outs.append(torch.matmul(probs, vb).permute(0, 2, 1, 3))   # [N, L, H, D]
    return torch.stack(outs, 0)


class EvoformerFusedAttention:
    """Kept for API parity (``EvoformerFusedAttention.apply(q, k, v, bias1, bias2)``); autograd comes from the fused
    ops underneath, so this is a thin callable rather than a hand-written ``autograd.Function``."""

    @staticmethod
    def apply(q, k, v, bias1=None, bias2=None):
        return _evoformer_core(q.contiguous(), k.contiguous(), v.contiguous(), bias1, bias2)


def DS4Sci_EvoformerAttention(Q: torch.Tensor, K: torch.Tensor, V: torch.Tensor,
                              biases: List[Optional[torch.Tensor]]) -> torch.Tensor:
    assert len(biases) <= 2
    biases = list(biases) + [None] * (2 - len(biases))
    squeeze = False
    if Q.dim() == 4:      # [N, L, H, D] -> add the batch dim
        Q, K, V = Q.unsqueeze(0), K.unsqueeze(0), V.unsqueeze(0)
        biases = [b.unsqueeze(0) if b is not None and b.dim() == 4 else b for b in biases]
        squeeze = True
    if biases[0] is not None:
        assert biases[0].shape == (Q.shape[0], Q.shape[1], 1, 1, Q.shape[2]), "bias1 shape is incorrect"
    if biases[1] is not None:
        assert biases[1].shape == (Q.shape[0], 1, Q.shape[3], Q.shape[2], Q.shape[2]), "bias2 shape is incorrect"
    out = EvoformerFusedAttention.apply(Q, K, V, biases[0], biases[1])
    return out.squeeze(0) if squeeze else out
