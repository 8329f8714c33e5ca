"""GPT-2 (BASELINE.json config 1: "GPT-2 small ZeRO-1 on CPU/gloo world_size=2").

Pre-LN decoder with learned positions, biased projections and tanh-GELU MLP; same fused-op structure
as ``models/llama.py`` (LayerNorm with fused residual add, bias+GELU kernel, tcgen05 GEMMs and flash
attention on CUDA, PyTorch composites on CPU)."""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.nn as nn

from ..ops import functional as F


@dataclass
class GPT2Config:
    vocab_size: int = 50304          # 50257 padded to a multiple of 64
    n_positions: int = 1024
    n_embd: int = 768
    n_layer: int = 12
    n_head: int = 12
    layer_norm_epsilon: float = 1e-5
    initializer_range: float = 0.02
    activation_checkpointing: bool = False

    def num_parameters(self):
        h = self.n_embd
        return self.n_layer * (12 * h * h + 13 * h) + (self.vocab_size + self.n_positions) * h + 2 * h


PRESETS = {
    "gpt2-small": dict(),
    "gpt2-medium": dict(n_embd=1024, n_layer=24, n_head=16),
    "gpt2-large": dict(n_embd=1280, n_layer=36, n_head=20),
    "gpt2-xl": dict(n_embd=1600, n_layer=48, n_head=25),
    "gpt2-tiny": dict(vocab_size=512, n_positions=128, n_embd=64, n_layer=2, n_head=2),
}


def gpt2_config(name="gpt2-small", **kw):
    d = dict(PRESETS[name])
    d.update(kw)
    return GPT2Config(**d)


class GPT2Block(nn.Module):
    def __init__(self, cfg: GPT2Config, device=None, dtype=None):
        super().__init__()
        kw = dict(device=device, dtype=dtype)
        h = cfg.n_embd
        self.cfg = cfg
        self.ln_1_w = nn.Parameter(torch.ones(h, **kw))
        self.ln_1_b = nn.Parameter(torch.zeros(h, **kw))
        self.c_attn_w = nn.Parameter(torch.empty(3 * h, h, **kw))
        self.c_attn_b = nn.Parameter(torch.zeros(3 * h, **kw))
        self.c_proj_w = nn.Parameter(torch.empty(h, h, **kw))
        self.c_proj_b = nn.Parameter(torch.zeros(h, **kw))
        self.ln_2_w = nn.Parameter(torch.ones(h, **kw))
        self.ln_2_b = nn.Parameter(torch.zeros(h, **kw))
        self.c_fc_w = nn.Parameter(torch.empty(4 * h, h, **kw))
        self.c_fc_b = nn.Parameter(torch.zeros(4 * h, **kw))
        self.c_proj2_w = nn.Parameter(torch.empty(h, 4 * h, **kw))
        self.c_proj2_b = nn.Parameter(torch.zeros(h, **kw))
        for w in (self.c_attn_w, self.c_proj_w, self.c_fc_w, self.c_proj2_w):
            nn.init.normal_(w, 0.0, cfg.initializer_range)

    def forward(self, h, residual):
        cfg = self.cfg
        B, S, H = h.shape
        nh, d = cfg.n_head, H // cfg.n_head
        x, residual = F.layer_norm(h, self.ln_1_w, self.ln_1_b, cfg.layer_norm_epsilon, residual)
        qkv = F.linear(x, self.c_attn_w, self.c_attn_b)
        if h.is_cuda and d not in (64, 128):
            raise RuntimeError("head_dim must be 64 or 128 on the CUDA path")
        o = F.flash_attention_qkv(qkv, nh, nh, d, causal=True)
        h = F.linear(o.reshape(B, S, H), self.c_proj_w, self.c_proj_b)
        x, residual = F.layer_norm(h, self.ln_2_w, self.ln_2_b, cfg.layer_norm_epsilon, residual)
        a = F.bias_act(F.linear(x, self.c_fc_w), self.c_fc_b, "gelu")
        h = F.linear(a, self.c_proj2_w, self.c_proj2_b)
        return h, residual


class GPT2LMHeadModel(nn.Module):
    def __init__(self, cfg: GPT2Config, device=None, dtype=None):
        super().__init__()
        self.cfg = self.config = cfg
        kw = dict(device=device, dtype=dtype)
        self.wte = nn.Parameter(torch.empty(cfg.vocab_size, cfg.n_embd, **kw))
        self.wpe = nn.Parameter(torch.empty(cfg.n_positions, cfg.n_embd, **kw))
        self.h = nn.ModuleList([GPT2Block(cfg, **kw) for _ in range(cfg.n_layer)])
        self.ln_f_w = nn.Parameter(torch.ones(cfg.n_embd, **kw))
        self.ln_f_b = nn.Parameter(torch.zeros(cfg.n_embd, **kw))
        nn.init.normal_(self.wte, 0.0, cfg.initializer_range)
        nn.init.normal_(self.wpe, 0.0, cfg.initializer_range)

    def forward(self, input_ids, labels=None):
        B, S = input_ids.shape
        pos = torch.arange(S, device=input_ids.device).unsqueeze(0).expand(B, S)
        h = F.embedding(input_ids, self.wte) + F.embedding(pos, self.wpe)
        residual = None
        ckpt = self.cfg.activation_checkpointing and self.training and torch.is_grad_enabled()
        for blk in self.h:
            if ckpt:
                from ..runtime.activation_checkpointing import checkpointing
                h, residual = checkpointing.checkpoint(blk, h, residual)
            else:
                h, residual = blk(h, residual)
        x, _ = F.layer_norm(h, self.ln_f_w, self.ln_f_b, self.cfg.layer_norm_epsilon, residual)
        logits = F.linear(x, self.wte)         # tied LM head
        if labels is None:
            return logits
        tgt = torch.full_like(labels, -100)
        tgt[:, :-1] = labels[:, 1:]
        return F.cross_entropy(logits, tgt, ignore_index=-100)
