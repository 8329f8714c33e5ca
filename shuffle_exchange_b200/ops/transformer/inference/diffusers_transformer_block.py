"""Fused BasicTransformerBlock of diffusers UNets (reference ``diffusers_transformer_block.py:33-108``): three
LayerNorms, self attention, cross attention and a GEGLU feed-forward, with residual adds folded into the norm kernels
(``layer_norm(x + residual)`` returns both the normed value and the sum)."""
from __future__ import annotations

import torch
import torch.nn as nn

from ... import functional as F
from .diffusers_2d_transformer import Diffusers2DTransformerConfig


class DeepSpeedDiffusersTransformerBlock(nn.Module):

    def __init__(self, equivalent_module: nn.Module, config: Diffusers2DTransformerConfig):
        super().__init__()
        self.config = config
        m = equivalent_module
        own = lambda t: nn.Parameter(t.data, requires_grad=False)   # noqa: E731
        ff_in = m.ff.net[0].proj
        self.ff1_w, self.ff1_b = own(ff_in.weight), own(ff_in.bias)
        self.ff2_w, self.ff2_b = own(m.ff.net[2].weight), own(m.ff.net[2].bias)
        self.norm1_g, self.norm1_b, self.norm1_eps = own(m.norm1.weight), own(m.norm1.bias), m.norm1.eps
        self.norm2_g, self.norm2_b, self.norm2_eps = own(m.norm2.weight), own(m.norm2.bias), m.norm2.eps
        self.norm3_g, self.norm3_b, self.norm3_eps = own(m.norm3.weight), own(m.norm3.bias), m.norm3.eps
        self.attn_1, self.attn_2 = m.attn1, m.attn2
        zero = lambda a: nn.Parameter(torch.zeros_like(self.norm2_g), requires_grad=False)   # noqa: E731
        self.attn_1_bias = getattr(self.attn_1, "attn_ob", None)
        if self.attn_1_bias is not None and hasattr(self.attn_1, "do_out_bias"):
            self.attn_1.do_out_bias = False
        else:
            self.attn_1_bias = zero(None)
        self.attn_2_bias = getattr(self.attn_2, "attn_ob", None)
        if self.attn_2_bias is not None and hasattr(self.attn_2, "do_out_bias"):
            self.attn_2.do_out_bias = False
        else:
            self.attn_2_bias = zero(None)

    def forward(self, hidden_states, context=None, timestep=None, **kwargs):
        if "encoder_hidden_states" in kwargs and kwargs["encoder_hidden_states"] is not None:
            context = kwargs["encoder_hidden_states"]
        h1 = F.layer_norm(hidden_states, self.norm1_g, self.norm1_b, self.norm1_eps)[0]
        a1 = self.attn_1(h1)
        h2, res1 = F.layer_norm(a1 + self.attn_1_bias, self.norm2_g, self.norm2_b, self.norm2_eps,
                                residual=hidden_states)
        a2 = self.attn_2(h2, context) if context is not None else self.attn_2(h2)
        h3, res2 = F.layer_norm(a2 + self.attn_2_bias, self.norm3_g, self.norm3_b, self.norm3_eps, residual=res1)
        gg = F.linear(h3, self.ff1_w, self.ff1_b)
        val, gate = gg.chunk(2, dim=-1)
        inter = val * torch.nn.functional.gelu(gate)
        return F.linear(inter, self.ff2_w, self.ff2_b) + res2
