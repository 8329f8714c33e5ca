"""MoE transformer layer for v1 inference.

Reference: ``ops/transformer/inference/moe_inference.py`` (``DeepSpeedMoEInferenceConfig`` :18-106,
``DeepSpeedMoEMLP`` :134-168, ``DeepSpeedMoEInference`` :171-360): attention -> norm -> top-k gate -> expert-parallel
all-to-all -> local experts -> all-to-all -> weighted combine (+ optional residual-MoE branch mixed by a learned
2-way softmax coefficient).

The reference dispatches with a dense one-hot ``einsum('sec,sm->ecm')`` into fixed-capacity buffers.  Here tokens are
routed by the framework's dropless kernels (``sxb_top_k_gating`` / ``sxb_moe_scatter`` / ``sxb_moe_gather``): rows are
grouped by expert with exact counts, exchanged with a split-size all-to-all, and every expert runs one GEMM pair over
just its rows — no capacity padding, no dropped tokens."""
from __future__ import annotations

import json
from typing import List

import torch
import torch.nn as nn

from .... import comm as dist
from ... import functional as F
from ... import moe_ops
from .config import DeepSpeedInferenceConfig
from .ds_attention import DeepSpeedSelfAttention, _device
from .op_binding import BiasResidualOp, LayerNormOp, MoEResMatmulOp, VectorMatMulOp


class DeepSpeedMoEInferenceConfig(DeepSpeedInferenceConfig):

    def __init__(self, hidden_size=-1, intermediate_size=-1, heads=-1, num_hidden_layers=-1, layer_norm_eps=1e-12,
                 local_rank=-1, mp_size=1, fp16=False, bf16=False, q_int8=False, pre_layer_norm=True,
                 stochastic_mode=False, scale_attention=True, triangular_masking=True, local_attention=False,
                 window_size=256, return_tuple=True, moe_experts=1, global_experts=1, k=1, capacity_factor=1.,
                 eval_capacity_factor=1., min_capacity=1, noisy_gate_policy=None, drop_tokens=True, use_rts=False,
                 mlp_type='standard', scale_attn_by_inverse_layer_idx=False):
        dtype = torch.int8 if q_int8 else (torch.bfloat16 if bf16 else (torch.half if fp16 else torch.float))
        super().__init__(hidden_size, intermediate_size if intermediate_size > 0 else 4 * hidden_size, heads,
                         num_hidden_layers, layer_norm_eps, local_rank, mp_size, dtype, pre_layer_norm,
                         stochastic_mode=stochastic_mode, scale_attention=scale_attention,
                         triangular_masking=triangular_masking, local_attention=local_attention,
                         window_size=window_size, return_tuple=return_tuple)
        self.fp16, self.bf16, self.q_int8 = fp16, bf16, q_int8
        self.moe_experts = moe_experts
        self.k = k
        self.capacity_factor = capacity_factor
        self.eval_capacity_factor = eval_capacity_factor
        self.min_capacity = min_capacity
        self.noisy_gate_policy = noisy_gate_policy
        self.drop_tokens = drop_tokens
        self.use_rts = use_rts
        self.global_experts = global_experts
        self.mlp_type = mlp_type
        self.scale_attn_by_inverse_layer_idx = scale_attn_by_inverse_layer_idx

    @classmethod
    def from_dict(cls, json_object):
        config = cls()
        for key, value in json_object.items():
            config.__dict__[key] = value
        return config

    @classmethod
    def from_json_file(cls, json_file):
        with open(json_file, "r", encoding="utf-8") as reader:
            return cls.from_dict(json.loads(reader.read()))


class DeepSpeedMoEMLP(nn.Module):
    """One expert: ``gelu(x W1 + b1) W2 + b2`` with v1 ``[in, out]`` weights."""

    def __init__(self, config, q_scales=None, q_groups=1, merge_count=1, mlp_extra_grouping=False, mp_group=None):
        super().__init__()
        self.config = config
        dt = torch.half if config.dtype == torch.int8 else config.dtype
        dev = _device()
        inter = config.intermediate_size // (1 if mp_group is None or not dist.is_initialized()
                                             else dist.get_world_size(group=mp_group))
        mk = lambda *s: nn.Parameter(torch.empty(*s, dtype=dt, device=dev), requires_grad=False)   # noqa: E731
        self.attn_nw, self.attn_nb = mk(config.hidden_size), mk(config.hidden_size)
        self.inter_w, self.inter_b = mk(config.hidden_size, inter), mk(inter)
        self.output_w, self.output_b = mk(inter, config.hidden_size), mk(config.hidden_size)
        self.mp_group = mp_group

    def forward(self, input, async_op=False):
        h = F.bias_act(F.linear(input, self.inter_w.t()), self.inter_b, "gelu")
        out = F.linear(h, self.output_w.t())
        if self.mp_group is not None and dist.is_initialized() and dist.get_world_size(group=self.mp_group) > 1:
            dist.all_reduce(out, group=self.mp_group)
        return out + self.output_b


class DeepSpeedMoEInference(nn.Module):
    layer_id = 0

    def __init__(self, config, mp_group=None, ep_group=None, expert_mp_group=None, quantize_scales=None,
                 quantize_groups=1, merge_count=1, mlp_extra_grouping=False):
        super().__init__()
        self.config = config
        self.config.layer_id = DeepSpeedMoEInference.layer_id
        DeepSpeedMoEInference.layer_id += 1
        dt = torch.half if config.dtype == torch.int8 else config.dtype
        dev = _device()
        self.attention = DeepSpeedSelfAttention(config, mp_group, quantize_scales, quantize_groups, merge_count)
        mk = lambda *s: nn.Parameter(torch.empty(*s, dtype=dt, device=dev), requires_grad=False)   # noqa: E731
        self.attn_nw, self.attn_nb = mk(config.hidden_size), mk(config.hidden_size)
        self.norm_w, self.norm_b = mk(config.hidden_size), mk(config.hidden_size)
        if config.mlp_type == "residual":
            self.res_mlp = DeepSpeedMoEMLP(config, quantize_scales, quantize_groups, merge_count, mlp_extra_grouping,
                                           mp_group)
            self.res_coef = mk(config.hidden_size, 2)
            self.vector_matmul_func = VectorMatMulOp(config)
        config.mp_size = 1
        self.mlp = nn.ModuleList(DeepSpeedMoEMLP(config, quantize_scales, quantize_groups, merge_count,
                                                 mlp_extra_grouping, expert_mp_group)
                                 for _ in range(config.moe_experts))
        self.moe_gate_w = mk(config.global_experts, config.hidden_size)       # router, [experts, hidden]
        self.ep_group, self.mp_group, self.expert_mp_group = ep_group, mp_group, expert_mp_group
        self.bias_residual_func = BiasResidualOp(config)
        self.ds_layernorm = LayerNormOp(config)
        self.moe_res_matmul = MoEResMatmulOp(config)

    # ------------------------------------------------------------------------------------------ routing
    def _ep(self) -> int:
        if self.ep_group is None or not dist.is_initialized():
            return 1
        return dist.get_world_size(group=self.ep_group)

    def _run_local(self, rows: torch.Tensor, ends: List[int]) -> torch.Tensor:
        out = torch.empty_like(rows)
        lo = 0
        for e, hi in enumerate(ends):
            if hi > lo:
                out[lo:hi] = self.mlp[e](rows[lo:hi])
            lo = hi
        return out

    def moe_forward(self, x2d: torch.Tensor) -> torch.Tensor:
        cfg = self.config
        logits = torch.nn.functional.linear(x2d, self.moe_gate_w)
        scores, assign, offsets, counts = moe_ops.top_k_gating(logits, cfg.k, renormalize=cfg.k > 1)
        moe_in, cumsum, slots = moe_ops.moe_scatter(x2d, assign, offsets, counts)
        ep, L = self._ep(), cfg.moe_experts
        if ep == 1:
            rows = self._run_local(moe_in, cumsum.tolist())
        else:
            send_counts = counts.view(ep, L)
            recv_counts = torch.empty_like(send_counts)
            dist.all_to_all_single(recv_counts, send_counts.contiguous(), group=self.ep_group)
            both = torch.stack([send_counts, recv_counts]).cpu()
            send_splits, recv_splits = both[0].sum(1).tolist(), both[1].sum(1).tolist()
            recv = moe_in.new_empty(sum(recv_splits), moe_in.shape[1])
            dist.all_to_all_single(recv, moe_in, recv_splits, send_splits, group=self.ep_group)
            rc = both[1]
            starts = torch.cumsum(rc.reshape(-1), 0) - rc.reshape(-1)
            idx = torch.cat([torch.arange(int(starts[s * L + e]), int(starts[s * L + e] + rc[s, e]))
                             for e in range(L) for s in range(ep)]).to(recv.device)
            grouped = recv.index_select(0, idx)
            out_g = self._run_local(grouped, torch.cumsum(rc.sum(0), 0).tolist())
            out_local = torch.empty_like(out_g).index_copy(0, idx, out_g)
            rows = moe_in.new_empty(moe_in.shape)
            dist.all_to_all_single(rows, out_local, send_splits, recv_splits, group=self.ep_group)
        return moe_ops.moe_gather(rows, scores.to(torch.float32), slots)

    def res_coef_func(self, inp, async_op=False):
        return torch.softmax(self.vector_matmul_func(inp, self.res_coef).float(), dim=-1).to(inp.dtype)

    # ------------------------------------------------------------------------------------------ layer
    def forward(self, input, input_mask=None, attention_mask=None, head_mask=None, layer_past=None,
                get_key_value=False, get_present=False, encoder_output=None, enc_dec_attn_mask=None,
                encoder_hidden_states=None, encoder_attention_mask=None, use_cache=False, output_attentions=False):
        get_present = get_present or get_key_value or use_cache
        input_mask = input_mask if attention_mask is None else attention_mask
        input_type = input.dtype
        if self.config.dtype in (torch.float16, torch.int8) and input_type == torch.float:
            input = input.half()
        with torch.no_grad():
            att = self.attention(input, input_mask, head_mask, layer_past, get_present, encoder_hidden_states,
                                 encoder_attention_mask, output_attentions, self.norm_w, self.norm_b)
            attention_output, presents = att[0], (att[1], att[2])
            residual_add = attention_output + self.attention.attn_ob + input
            normed = self.ds_layernorm(residual_add, self.attn_nw, self.attn_nb, self.config.epsilon)
            if self.config.mlp_type == "residual":
                res_mlp_out = self.res_mlp(normed)
                res_coef_out = self.res_coef_func(normed)
            x = normed
            if self.expert_mp_group is not None and dist.is_initialized() \
                    and dist.get_world_size(group=self.expert_mp_group) > 1:
                w = dist.get_world_size(group=self.expert_mp_group)
                buf = torch.empty(w * x.numel(), dtype=x.dtype, device=x.device)
                dist.all_gather_into_tensor(buf, x.contiguous(), group=self.expert_mp_group)
                x = buf.view(-1, *x.shape[1:])
            out = self.moe_forward(x.reshape(-1, self.config.hidden_size)).reshape(x.shape)
            if x is not normed:
                r = dist.get_rank(group=self.expert_mp_group)
                out = out.split(out.shape[0] // dist.get_world_size(group=self.expert_mp_group), dim=0)[r]
            if self.config.mlp_type == "residual":
                out = self.moe_res_matmul(res_mlp_out, res_coef_out, out)
            output = out + residual_add
            if not self.config.pre_layer_norm:
                output = self.ds_layernorm(output, self.norm_w, self.norm_b, self.config.epsilon)
            if input_type != output.dtype:
                output = output.to(input_type)
        if get_present:
            output = (output, presents)
        if self.config.return_tuple:
            return output if type(output) is tuple else (output,)
        return output
