"""Mixtral-style sparse decoder: Llama attention + a top-k mixture of SwiGLU experts per layer.

The reference ships Mixtral only as an inference-v2 model (``inference/v2/model_implementations/mixtral``) and trains
MoE models through ``deepspeed.moe.layer.MoE``; the headline benchmark names "Mixtral EP".  This module is the
training model: attention path identical to :mod:`models.llama` (fused QKV, RoPE, flash attention), MLP replaced by
:class:`~shuffle_exchange_b200.moe.dropless.DroplessMoE` with the experts sharded over the expert-parallel group.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import torch
import torch.nn as nn

from ..moe.dropless import DroplessMoE
from ..ops import functional as F
from .llama import LlamaConfig


@dataclass
class MixtralConfig(LlamaConfig):
    vocab_size: int = 32000
    rope_theta: float = 1000000.0
    max_position_embeddings: int = 32768
    num_local_experts: int = 8
    num_experts_per_tok: int = 2
    router_aux_loss_coef: float = 0.02
    ep_size: int = 1
    fused_exchange: object = None        # None: symmetric-memory dispatch/combine when available; False: NCCL
    fp8_experts: bool = False            # e4m3 expert GEMMs in forward (tcgen05 kind::f8f6f4), bf16 backward
    ep_capacity_factor: float = 2.0          # 1.25 overflowed at 8 GPUs with a random-init router (overflow raises)

    def num_parameters(self) -> int:
        h, f, v, d = self.hidden_size, self.intermediate_size, self.vocab_size, self.head_dim
        attn = h * (self.num_attention_heads + 2 * self.num_key_value_heads) * d + h * h
        per_layer = attn + self.num_local_experts * 3 * h * f + self.num_local_experts * h + 2 * h
        return self.num_hidden_layers * per_layer + 2 * v * h + h

    def activation_bytes_per_layer(self, batch: int, seq_len: int, elem: int = 2) -> int:
        """Dense part as in Llama (residuals, normed inputs, qkv, attention output) plus the routed rows: dispatched
        copy, gate|up, SwiGLU product and expert output for ``k`` rows per token (worst case all on this rank's
        experts under expert parallelism: tokens arrive from every rank, ~k*t rows on average)."""
        t = batch * seq_len
        h, f, k = self.hidden_size, self.intermediate_size, self.num_experts_per_tok
        qkv = (self.num_attention_heads + 2 * self.num_key_value_heads) * self.head_dim
        # the fused exchange sizes every routed-row tensor for the receive-buffer capacity (no host sync to learn the
        # exact row count), plus one 256-row alignment pad per local expert
        rows = t * k
        if self.fused_exchange is not False and self.ep_size > 1:
            rows = int(rows * self.ep_capacity_factor) + 256 * max(1, self.num_local_experts // self.ep_size)
        return t * elem * (5 * h + qkv) + rows * elem * (3 * h + 3 * f) + t * k * elem * h \
            + t * 4 * (2 + self.num_attention_heads)

    def flops_per_token(self, seq_len: int, causal: bool = True) -> float:
        h, f, d = self.hidden_size, self.intermediate_size, self.head_dim
        attn_mm = h * (self.num_attention_heads + 2 * self.num_key_value_heads) * d + h * h
        active = attn_mm + self.num_experts_per_tok * 3 * h * f + self.num_local_experts * h
        mm = self.num_hidden_layers * active + self.vocab_size * h
        attn = self.num_hidden_layers * 2 * 2 * self.num_attention_heads * d * seq_len * (0.5 if causal else 1.0)
        return 6.0 * mm + 3.0 * attn


MIXTRAL_PRESETS = {
    "mixtral-8x7b": dict(),
    "mixtral-tiny": dict(vocab_size=1024, hidden_size=256, intermediate_size=512, num_hidden_layers=2,
                         num_attention_heads=4, num_key_value_heads=2, max_position_embeddings=512,
                         num_local_experts=4, num_experts_per_tok=2),
}


def mixtral_config(name: str, **overrides) -> MixtralConfig:
    kw = dict(MIXTRAL_PRESETS[name])
    kw.update(overrides)
    return MixtralConfig(**kw)


class MixtralDecoderLayer(nn.Module):
    def __init__(self, cfg: MixtralConfig, layer_idx: int, device=None, dtype=None):
        super().__init__()
        kw = dict(device=device, dtype=dtype)
        h, d = cfg.hidden_size, cfg.head_dim
        self.cfg, self.layer_idx = cfg, layer_idx
        self.n_q, self.n_kv, self.d = cfg.num_attention_heads, cfg.num_key_value_heads, d
        self.input_layernorm = nn.Parameter(torch.ones(h, **kw))
        self.wqkv = nn.Parameter(torch.empty((self.n_q + 2 * self.n_kv) * d, h, **kw))
        self.wo = nn.Parameter(torch.empty(h, self.n_q * d, **kw))
        self.post_attention_layernorm = nn.Parameter(torch.ones(h, **kw))
        self.moe = DroplessMoE(h, cfg.intermediate_size, cfg.num_local_experts, ep_size=cfg.ep_size,
                               k=cfg.num_experts_per_tok, aux_loss_coef=cfg.router_aux_loss_coef,
                               fused_exchange=cfg.fused_exchange, fp8_experts=cfg.fp8_experts,
                               ep_capacity_factor=cfg.ep_capacity_factor, **kw)

    def reset_parameters(self, std: float):
        for w in (self.wqkv, self.wo):
            if w.device.type != "meta":
                nn.init.normal_(w, 0.0, std)
        self.moe.reset_parameters(std)

    def forward(self, h, residual, positions, cos, sin):
        B, S, _ = h.shape
        eps = self.cfg.rms_norm_eps
        x, residual = F.rms_norm(h, self.input_layernorm, eps, residual)
        qkv = F.linear(x, self.wqkv)
        qkv = F.rope(qkv, positions, cos, sin, self.n_q + self.n_kv, self.d)
        o = F.flash_attention_qkv(qkv, self.n_q, self.n_kv, self.d, causal=True)
        h = F.linear(o.reshape(B, S, self.n_q * self.d), self.wo)
        x, residual = F.rms_norm(h, self.post_attention_layernorm, eps, residual)
        h, l_aux, _ = self.moe(x)
        return h, residual, l_aux


class MixtralModel(nn.Module):
    def __init__(self, cfg: MixtralConfig, device=None, dtype=None):
        super().__init__()
        self.cfg = cfg
        kw = dict(device=device, dtype=dtype)
        self.embed_tokens = nn.Parameter(torch.empty(cfg.vocab_size, cfg.hidden_size, **kw))
        self.layers = nn.ModuleList([MixtralDecoderLayer(cfg, i, **kw) for i in range(cfg.num_hidden_layers)])
        self.norm = nn.Parameter(torch.ones(cfg.hidden_size, **kw))

    def forward(self, input_ids, position_ids=None):
        B, S = input_ids.shape
        cfg, dev = self.cfg, input_ids.device
        positions = position_ids.reshape(-1) if position_ids is not None else \
            torch.arange(S, device=dev, dtype=torch.long).repeat(B)
        cos, sin = F.rope_tables(max(cfg.max_position_embeddings, S), cfg.head_dim, cfg.rope_theta, dev)
        h = F.embedding(input_ids, self.embed_tokens)
        residual, aux = None, 0.0
        ckpt = cfg.activation_checkpointing and self.training and torch.is_grad_enabled()
        n_ckpt = len(self.layers) if cfg.checkpoint_layers < 0 else cfg.checkpoint_layers
        for i, layer in enumerate(self.layers):
            if ckpt and i < n_ckpt:
                from ..runtime.activation_checkpointing import checkpointing
                h, residual, l_aux = checkpointing.checkpoint(layer, h, residual, positions, cos, sin)
            else:
                h, residual, l_aux = layer(h, residual, positions, cos, sin)
            aux = aux + l_aux
        x, _ = F.rms_norm(h, self.norm, cfg.rms_norm_eps, residual)
        return x, aux


class MixtralForCausalLM(nn.Module):
    def __init__(self, cfg: MixtralConfig, device=None, dtype=None):
        super().__init__()
        self.cfg = self.config = cfg
        self.model = MixtralModel(cfg, device=device, dtype=dtype)
        self.lm_head = nn.Parameter(torch.empty(cfg.vocab_size, cfg.hidden_size, device=device, dtype=dtype))
        self.reset_parameters()

    def reset_parameters(self):
        std = self.cfg.initializer_range
        if self.lm_head.device.type != "meta":
            nn.init.normal_(self.model.embed_tokens, 0.0, std)
            nn.init.normal_(self.lm_head, 0.0, std)
        for l in self.model.layers:
            l.reset_parameters(std)

    def head_weight(self):
        return self.lm_head

    def forward(self, input_ids, labels=None, position_ids=None, shift_labels: bool = True):
        x, aux = self.model(input_ids, position_ids)
        logits = F.linear(x, self.lm_head)
        if labels is None:
            return logits
        if shift_labels:
            tgt = torch.full_like(labels, -100)
            tgt[:, :-1] = labels[:, 1:]
        else:
            tgt = labels
        loss = F.cross_entropy(logits, tgt, ignore_index=-100)
        if self.cfg.router_aux_loss_coef > 0:
            loss = loss + self.cfg.router_aux_loss_coef * aux
        return loss
