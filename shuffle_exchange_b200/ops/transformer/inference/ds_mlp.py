"""MLP block of the v1 inference transformer (reference ``ops/transformer/inference/ds_mlp.py:15-125``): parameters
``attn_nw/attn_nb`` (post-attention norm), ``inter_w [hidden, inter*(1|2)/mp]``, ``inter_b``, ``output_w [inter/mp,
hidden]``, ``output_b``; ``forward(input, residual, residual_norm, bias)``; all-reduce of the output over the TP
group."""
from __future__ import annotations

import math

import torch
import torch.nn as nn

from .... import comm as dist
from ....utils.types import GATED_ACTIVATION_TYPES
from .ds_attention import _device
from .op_binding import GELUGemmOp, MLPGemmOp, ResidualAddOp, VectorMatMulOp


class DeepSpeedMLP(nn.Module):
    _inter_w_buffers = []

    def __init__(self, config, mp_group=None, q_scales=None, q_groups=1, merge_count=1, mlp_extra_grouping=False):
        super().__init__()
        self.config = config
        dt = torch.half if config.dtype == torch.int8 else config.dtype
        dev = _device()
        factor = 2 if config.mlp_act_func_type in GATED_ACTIVATION_TYPES else 1
        config.intermediate_size = config.intermediate_size if config.intermediate_size > 0 else 4 * config.hidden_size
        self.intm_w_sz_per_partition = config.intermediate_size * factor // config.mp_size
        self.intm_o_sz_per_partition = config.intermediate_size // config.mp_size
        if config.set_empty_params:
            for n in ("attn_nw", "attn_nb", "inter_w", "inter_b", "inter_up_w", "inter_up_b", "inter_gate_w",
                      "inter_gate_b", "output_w", "output_b"):
                setattr(self, n, None)
        else:
            mk = lambda *s: nn.Parameter(torch.empty(*s, dtype=dt, device=dev), requires_grad=False)   # noqa: E731
            self.attn_nw, self.attn_nb = mk(config.hidden_size), mk(config.hidden_size)
            self.inter_w, self.inter_b = mk(config.hidden_size, self.intm_w_sz_per_partition), \
                mk(self.intm_w_sz_per_partition)
            self.output_w, self.output_b = mk(self.intm_o_sz_per_partition, config.hidden_size), mk(config.hidden_size)
        self.q_scales = q_scales
        self.q_groups = q_groups * 2 if mlp_extra_grouping else q_groups
        self.merge_count = int(math.log2(merge_count))
        self.mp_group = mp_group
        self.mlp_gemm_func = MLPGemmOp(config)
        self.vector_matmul_func = VectorMatMulOp(config)
        self.fused_gemm_gelu = GELUGemmOp(config)
        self.residual_add_func = ResidualAddOp(config)

    def _merge_inter_w(self):
        dim = 0 if self.config.transposed_mode else 1
        w = torch.cat([self.inter_up_w, self.inter_gate_w], dim=dim)
        b = torch.cat([self.inter_up_b, self.inter_gate_b]) if self.inter_up_b is not None else None
        return w, b

    def forward(self, input, residual, residual_norm, bias):
        inter_w, inter_b = (self.inter_w, self.inter_b) if self.inter_w is not None else self._merge_inter_w()
        residual_add = None
        if self.attn_nw is None:
            output = self.fused_gemm_gelu(input=residual_norm, weight=inter_w, bias=inter_b, weight_out=self.output_w)
        else:
            output, residual_add = self.mlp_gemm_func(input=input, residual=residual, weight_interm=inter_w,
                                                      weight_out=self.output_w, input_bias=bias, bias=inter_b,
                                                      gamma=self.attn_nw, beta=self.attn_nb)
        tp = self.mp_group is not None and dist.is_initialized() and dist.get_world_size(group=self.mp_group) > 1
        if tp and not self.config.mlp_after_attn:
            # parallel attention + MLP block: the attention partial product was NOT reduced (ds_attention skips it),
            # so both partial sums travel in one all-reduce; residual and biases are added once afterwards
            output = output + input
            dist.all_reduce(output, group=self.mp_group)
            return self.residual_add_func(hidden_state=output, residual=residual, add_bias=bias is not None,
                                          attention_output=None, attention_bias=bias, final_bias=self.output_b,
                                          residual_add=residual_add)
        if tp:
            dist.all_reduce(output, group=self.mp_group)
        return self.residual_add_func(hidden_state=output, residual=residual, add_bias=bias is not None,
                                      attention_output=input,
                                      attention_bias=bias if bias is not None else self.output_b,
                                      final_bias=self.output_b, residual_add=residual_add)
