"""Self-attention block of the v1 inference transformer with an in-module KV cache.

Reference: ``ops/transformer/inference/ds_attention.py:15-290`` (``DeepSpeedSelfAttention`` + ``BloomSelfAttention``):
parameters ``attn_qkvw [hidden, qkv/mp]``, ``attn_qkvb``, ``attn_ow [hidden/mp, hidden]``, ``attn_ob`` (or split
``attn_{q,k,v}w`` when ``set_empty_params``), forward returns ``(output, key, value, context, normed_input)`` and
all-reduces the output projection over the tensor-parallel group."""
from __future__ import annotations

import math

import torch
import torch.nn as nn

from .... import comm as dist
from .op_binding import LinearOp, QKVGemmOp, SoftmaxContextOp, SoftmaxOp, VectorMatMulOp

minus_inf = -10000.0


def _device():
    return torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")


class DeepSpeedSelfAttention(nn.Module):
    num_layers = 0
    _qkv_buffers = []

    def __init__(self, config, mp_group=None, q_scales=None, q_groups=1, merge_count=1):
        super().__init__()
        self.config = config
        dt = torch.half if config.dtype == torch.int8 else config.dtype
        self.config.layer_id = DeepSpeedSelfAttention.num_layers
        DeepSpeedSelfAttention.num_layers += 1
        dev = _device()
        head_dim = config.hidden_size // config.heads
        if config.set_empty_params:
            for n in ("attn_qw", "attn_qb", "attn_kw", "attn_kb", "attn_vw", "attn_vb", "attn_qkvw", "attn_qkvb",
                      "attn_ow", "attn_ob"):
                setattr(self, n, None)
        else:
            qkv = (config.hidden_size // config.mp_size) * 3 if config.num_kv < 0 else \
                ((config.heads + config.num_kv * 2) // config.mp_size) * head_dim
            out_in = config.hidden_size // config.mp_size
            mk = lambda *s: nn.Parameter(torch.empty(*s, dtype=dt, device=dev), requires_grad=False)   # noqa: E731
            self.attn_qkvw, self.attn_qkvb = mk(config.hidden_size, qkv), mk(qkv)
            self.attn_ow, self.attn_ob = mk(out_in, config.hidden_size), mk(config.hidden_size)
        self.num_attention_heads_per_partition = config.heads // config.mp_size
        self.num_kv_partition = config.num_kv // config.mp_size if config.num_kv > 0 else -1
        self.hidden_size_per_partition = config.hidden_size // config.mp_size
        self.hidden_size_per_attention_head = head_dim
        self.mp_group = mp_group
        self.q_scales, self.q_groups, self.merge_count = q_scales, q_groups, int(math.log2(merge_count))
        self.norm_factor = math.sqrt(head_dim)
        if not config.use_mup:
            self.norm_factor = math.sqrt(self.norm_factor)
        if not getattr(config, "scale_attention", True):
            self.norm_factor = 1.0                      # GPT-Neo: raw dot products
        if config.scale_attn_by_inverse_layer_idx:
            self.norm_factor *= math.sqrt(config.layer_id + 1)
        self.qkv_func = QKVGemmOp(config)
        self.score_context_func = SoftmaxContextOp(config)
        self.linear_func = LinearOp(config)
        self.vector_matmul_func = VectorMatMulOp(config)

    def _merge_qkv(self):
        """Split q/k/v weights (``set_empty_params`` containers) -> one fused ``[hidden, 3*hidden/mp]`` weight."""
        w = torch.cat([self.attn_qw, self.attn_kw, self.attn_vw], dim=0 if self.config.transposed_mode else 1)
        b = None
        if self.attn_qb is not None:
            b = torch.cat([self.attn_qb, self.attn_kb, self.attn_vb], dim=0)
        return w, b

    def compute_attention(self, qkv_out, input_mask, layer_past, alibi, is_prompt, token_idx, position_ids):
        if isinstance(qkv_out, (list, tuple)):
            qkv_out = qkv_out[0]
        no_masking = input_mask is None or not self.config.triangular_masking
        mask = input_mask
        if mask is not None and self.config.invert_mask and mask.dtype not in (torch.float16, torch.bfloat16,
                                                                                 torch.float32):
            mask = (1 - mask.to(qkv_out.dtype)) * minus_inf
        if mask is not None and mask.dim() == 2:
            mask = mask[:, None, None, :]
        ctx, k, v = self.score_context_func(
            query_key_value=qkv_out, attn_mask=mask, heads=self.num_attention_heads_per_partition,
            num_kv=self.num_kv_partition, norm_factor=self.norm_factor, no_masking=False,
            layer_id=self.config.layer_id, num_layers=DeepSpeedSelfAttention.num_layers, alibi=alibi,
            is_prompt=is_prompt, token_idx=token_idx, position_ids=position_ids, layer_past=layer_past)
        return ctx, k, v

    def forward(self, input, input_mask, head_mask=None, layer_past=None, get_present=False,
                encoder_hidden_states=None, encoder_attention_mask=None, output_attentions=False, norm_w=None,
                norm_b=None, alibi=None, **kwargs):
        if self.attn_qkvw is None:
            qkvw, qkvb = self._merge_qkv()
        else:
            qkvw, qkvb = self.attn_qkvw, self.attn_qkvb
        if not self.config.pre_layer_norm:
            qkv_out = (self.linear_func(input=input, weight=qkvw, bias=qkvb, add_bias=qkvb is not None), input)
        else:
            qkv_out = self.qkv_func(input=input, weight=qkvw, bias=qkvb, gamma=norm_w, beta=norm_b)
        ctx, key, value = self.compute_attention(qkv_out=qkv_out, input_mask=input_mask, layer_past=layer_past,
                                                 alibi=alibi, is_prompt=input.shape[1] > 1,
                                                 token_idx=kwargs.get("token_idx"),
                                                 position_ids=kwargs.get("position_ids"))
        output = self.vector_matmul_func(input=ctx, weight=self.attn_ow)
        inp_norm = qkv_out[-1]
        if self.config.mlp_after_attn and self.mp_group is not None and dist.is_initialized() \
                and dist.get_world_size(group=self.mp_group) > 1:
            dist.all_reduce(output, group=self.mp_group)
        return output, key, value, ctx, inp_norm


class BloomSelfAttention(DeepSpeedSelfAttention):
    """BLOOM: per-head interleaved fused QKV ``[.., heads, 3, head_dim]`` and ALiBi biases."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self.softmax_func = SoftmaxOp(self.config)

    def compute_attention(self, qkv_out, input_mask, layer_past, alibi, is_prompt, token_idx, position_ids):
        if isinstance(qkv_out, (list, tuple)):
            qkv_out = qkv_out[0]
        B, S, W = qkv_out.shape
        H, D = self.num_attention_heads_per_partition, self.hidden_size_per_attention_head
        q, k, v = qkv_out.view(B, S, H, 3, D).unbind(3)
        regrouped = torch.cat([q.reshape(B, S, H * D), k.reshape(B, S, H * D), v.reshape(B, S, H * D)], dim=-1)
        return super().compute_attention(regrouped, input_mask, layer_past, alibi, is_prompt, token_idx, position_ids)
