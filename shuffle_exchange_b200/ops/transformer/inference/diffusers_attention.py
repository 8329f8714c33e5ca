"""Spatial self/cross attention of diffusers UNets (reference ``ops/transformer/inference/diffusers_attention.py``):
parameters ``attn_qkvw`` (self) or ``attn_qw/attn_kw/attn_vw`` (cross), ``attn_ow``, ``attn_ob``; input
``[B, HW, C]`` (or ``[B, C, H, W]``), optional ``context`` for cross attention."""
from __future__ import annotations

import math

import torch
import torch.nn as nn

from ... import functional as F
from .ds_attention import _device


class DeepSpeedDiffusersAttention(nn.Module):
    layer_id = 0

    def __init__(self, config):
        super().__init__()
        self.config = config
        self.config.layer_id = DeepSpeedDiffusersAttention.layer_id
        DeepSpeedDiffusersAttention.layer_id += 1
        dt = torch.half if config.dtype == torch.int8 else config.dtype
        dev = _device()
        h = config.hidden_size
        mk = lambda *s: nn.Parameter(torch.empty(*s, dtype=dt, device=dev), requires_grad=False)   # noqa: E731
        self.attn_qkvw = mk(h, 3 * h // config.mp_size)
        self.attn_kw, self.attn_vw, self.attn_qw = mk(h, h // config.mp_size), mk(h, h // config.mp_size), \
            mk(h, h // config.mp_size)
        self.attn_qkvb = mk(3 * h // config.mp_size)
        self.attn_ow, self.attn_ob = mk(h // config.mp_size, h), mk(h)
        self.do_out_bias = True
        self.num_attention_heads_per_partition = config.heads // config.mp_size
        self.hidden_size_per_partition = h // config.mp_size
        self.scale = 1.0 / math.sqrt(h // config.heads)

    def _attend(self, q, k, v):
        B, Sq, _ = q.shape
        H = self.num_attention_heads_per_partition
        D = q.shape[-1] // H
        q, k, v = (t.reshape(B, -1, H, D) for t in (q, k, v))
        if q.is_cuda and q.dtype == torch.bfloat16 and D in (64, 128) and k.shape[1] == Sq:
            out = F.flash_attention(q, k, v, causal=False, scale=self.scale)
        else:
            out = torch.nn.functional.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2),
                                                                   v.transpose(1, 2), scale=self.scale)
            out = out.transpose(1, 2)
        return out.reshape(B, Sq, H * D)

    def forward(self, input, context=None, input_mask=None):
        spatial = input.dim() == 4
        if spatial:
            n, c, hh, ww = input.shape
            input = input.reshape(n, c, hh * ww).transpose(1, 2)
        if context is None:
            qkv = F.linear(input, self.attn_qkvw.t(), self.attn_qkvb if self.attn_qkvb.numel() else None)
            q, k, v = qkv.chunk(3, dim=-1)
        else:
            q = F.linear(input, self.attn_qw.t())
            k = F.linear(context, self.attn_kw.t())
            v = F.linear(context, self.attn_vw.t())
        out = F.linear(self._attend(q, k, v), self.attn_ow.t(), self.attn_ob if self.do_out_bias else None)
        if spatial:
            out = out.transpose(1, 2).reshape(n, c, hh, ww)
        return out
