"""One transformer layer of the v1 inference path: attention block + MLP block over an in-layer KV cache.

Reference: ``model_implementations/transformers/ds_transformer.py:23-200`` — class-level layer counter and workspace,
``forward`` with the union of the keyword names the supported HF / Megatron layers use (``x`` / ``hidden_states``,
``input_mask`` / ``attention_mask`` / ``attn_mask``, ``layer_past`` / ``past_key_value``, ``get_present`` /
``get_key_value`` / ``use_cache`` …), return convention controlled by ``return_tuple`` / ``return_single_tuple``."""
from __future__ import annotations

import torch
import torch.nn as nn

from ... import comm as dist
from ...ops.transformer.inference.ds_attention import BloomSelfAttention, DeepSpeedSelfAttention, _device
from ...ops.transformer.inference.ds_mlp import DeepSpeedMLP
from ...ops.transformer.inference.op_binding import LayerNormOp, WorkspaceOp
from ...utils.logging import log_dist


class DeepSpeedTransformerInference(nn.Module):
    layer_id = 0
    workspace = None

    def __init__(self, config, mp_group=None, quantize_scales=None, quantize_groups=1, merge_count=1,
                 mlp_extra_grouping=False):
        super().__init__()
        self.config = config
        self.config.layer_id = DeepSpeedTransformerInference.layer_id
        DeepSpeedTransformerInference.layer_id += 1
        dt = torch.half if config.dtype == torch.int8 else config.dtype
        if DeepSpeedTransformerInference.layer_id == 1:
            log_dist(f"DeepSpeed-Inference config: {self.config.__dict__}", [0])
        attn_cls = BloomSelfAttention if config.bigscience_bloom else DeepSpeedSelfAttention
        self.attention = attn_cls(config, mp_group, quantize_scales, quantize_groups, merge_count)
        self.mlp = DeepSpeedMLP(config, mp_group, quantize_scales, quantize_groups, merge_count, mlp_extra_grouping)
        if config.set_empty_params:
            self.norm_w = self.norm_b = None
        else:
            dev = _device()
            self.norm_w = nn.Parameter(torch.empty(config.hidden_size, dtype=dt, device=dev), requires_grad=False)
            self.norm_b = nn.Parameter(torch.empty(config.hidden_size, dtype=dt, device=dev), requires_grad=False)
        self.layer_past = None
        self.layer_norm = LayerNormOp()
        if DeepSpeedTransformerInference.workspace is None:
            DeepSpeedTransformerInference.workspace = WorkspaceOp(config)
        self._should_allocate_workspace = True

    def allocate_workspace(self, size):
        if self.config.layer_id == 0 and self._should_allocate_workspace:
            DeepSpeedTransformerInference.workspace.allocate_workspace(
                self.config.hidden_size, self.config.heads, size[1], size[0], DeepSpeedTransformerInference.layer_id,
                self.config.mp_size, self.config.bigscience_bloom, dist.get_rank() if dist.is_initialized() else 0,
                self.config.max_out_tokens, self.config.min_out_tokens)
            self._should_allocate_workspace = False

    @classmethod
    def reset_cache(cls):
        if cls.workspace is not None:
            cls.workspace.reset_cache()

    def forward(self, input=None, input_mask=None, attention_mask=None, attn_mask=None, head_mask=None,
                layer_past=None, get_key_value=False, get_present=False, encoder_output=None,
                enc_dec_attn_mask=None, x=None, encoder_hidden_states=None, encoder_attention_mask=None,
                use_cache=False, alibi=None, output_attentions=False, layer_head_mask=None, past_key_value=None,
                **kwargs):
        if x is not None:
            input = x
        if "hidden_states" in kwargs:
            input = kwargs.pop("hidden_states")
        input_mask = (input_mask if attn_mask is None else attn_mask) if attention_mask is None else attention_mask
        extra_mask = None
        if isinstance(input, tuple):
            extra_mask, input = input[1], input[0]
        self.allocate_workspace(input.size())
        get_present = get_present or get_key_value or use_cache
        if input.shape[1] > 1:
            self.layer_past = None                      # a prompt starts a new sequence
        layer_past = layer_past if layer_past is not None else (past_key_value or self.layer_past)
        head_mask = layer_head_mask if layer_head_mask is not None else head_mask
        input_type = input.dtype
        if self.config.dtype in (torch.float16, torch.bfloat16, torch.int8) and input.dtype == torch.float:
            input = input.to(torch.half if self.config.dtype == torch.int8 else self.config.dtype)
        with torch.no_grad():
            attention_output, key, value, _ctx, inp_norm = self.attention(
                input, input_mask, head_mask, layer_past, get_present, encoder_hidden_states,
                encoder_attention_mask, output_attentions, self.norm_w, self.norm_b, alibi, **kwargs)
            presents = (key, value)
            self.layer_past = presents
            output = self.mlp(attention_output, input, inp_norm, self.attention.attn_ob)
            if not self.config.pre_layer_norm:
                output = self.layer_norm(output, self.norm_w, self.norm_b, self.config.epsilon)
            output = output.to(input_type)
        if get_present:
            output = (output, presents)
        if self.config.return_single_tuple:
            return (output,)
        if self.config.return_tuple:
            return output if type(output) is tuple else (output, extra_mask)
        return output
