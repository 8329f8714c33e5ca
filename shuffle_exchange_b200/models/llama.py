"""Llama-family decoder (Llama-2/3, Mistral-style GQA) built on the sm_100a kernels.

The reference does not ship model code for training — users bring a HF/Megatron module and the
engine wraps it.  This in-tree implementation is what ``bench.py`` trains (BASELINE.json config
"Llama-3 8B ZeRO-3 bf16") and what the tests use as the flagship model.  Design points:

* fused projections: one ``wqkv`` GEMM ([(Hq+2Hk)·D, H]) and one ``w_gate_up`` GEMM ([2F, H]) so the
  RoPE kernel rotates q|k in place inside the GEMM output and SwiGLU reads one contiguous buffer;
* the residual add is fused into RMSNorm (``rms_norm(h, w, eps, residual)`` returns the updated
  residual stream), so a block is GEMM → attention → GEMM → norm → GEMM → act → GEMM with no
  stand-alone elementwise kernels;
* every op is a hand-written kernel on CUDA (tcgen05 GEMM, tcgen05 flash attention, fused
  cross-entropy that overwrites the logits with their gradient) and a PyTorch composite on CPU;
* HF weight names are accepted by :meth:`LlamaForCausalLM.load_hf_state_dict`.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, asdict
from typing import Optional

import torch
import torch.nn as nn

from ..ops import functional as F


@dataclass
class LlamaConfig:
    vocab_size: int = 128256
    hidden_size: int = 4096
    intermediate_size: int = 14336
    num_hidden_layers: int = 32
    num_attention_heads: int = 32
    num_key_value_heads: int = 8
    max_position_embeddings: int = 8192
    rope_theta: float = 500000.0
    rms_norm_eps: float = 1e-5
    tie_word_embeddings: bool = False
    initializer_range: float = 0.02
    activation_checkpointing: bool = True
    # number of decoder layers (counted from layer 0) whose activations are recomputed in backward;
    # -1 = all of them.  Lets a run spend spare HBM on skipping recompute (see activation_bytes_per_layer).
    checkpoint_layers: int = -1

    @property
    def head_dim(self) -> int:
        return self.hidden_size // self.num_attention_heads

    def num_parameters(self) -> int:
        h, f, v = self.hidden_size, self.intermediate_size, self.vocab_size
        d = self.head_dim
        per_layer = h * (self.num_attention_heads + 2 * self.num_key_value_heads) * d + h * h + 3 * h * f + 2 * h
        return self.num_hidden_layers * per_layer + v * h * (1 if self.tie_word_embeddings else 2) + h

    def flops_per_token(self, seq_len: int, causal: bool = True) -> float:
        """Training FLOPs per token (fwd + bwd, no recompute): 6·N_matmul + attention."""
        h, f = self.hidden_size, self.intermediate_size
        d = self.head_dim
        mm = self.num_hidden_layers * (h * (self.num_attention_heads + 2 * self.num_key_value_heads) * d + h * h + 3 * h * f)
        mm += self.vocab_size * h
        attn = self.num_hidden_layers * 2 * 2 * self.num_attention_heads * d * seq_len * (0.5 if causal else 1.0)
        return 6.0 * mm + 3.0 * attn

    def activation_bytes_per_layer(self, batch: int, seq_len: int, elem: int = 2) -> int:
        """Bytes one decoder layer keeps for backward when it is NOT checkpointed: two residual streams, two
        normed inputs, qkv, attention output, gate|up and the SwiGLU product (plus fp32 rstd/lse rows)."""
        t = batch * seq_len
        h, f = self.hidden_size, self.intermediate_size
        qkv = (self.num_attention_heads + 2 * self.num_key_value_heads) * self.head_dim
        return t * elem * (5 * h + qkv + 3 * f) + t * 4 * (2 + self.num_attention_heads)

    to_dict = asdict


PRESETS = {
    "llama3-8b": dict(),
    "llama3-70b": dict(hidden_size=8192, intermediate_size=28672, num_hidden_layers=80, num_attention_heads=64,
                       num_key_value_heads=8),
    "llama2-7b": dict(vocab_size=32000, intermediate_size=11008, num_key_value_heads=32, rope_theta=10000.0,
                      max_position_embeddings=4096),
    "llama-tiny": dict(vocab_size=1024, hidden_size=256, intermediate_size=512, num_hidden_layers=2,
                       num_attention_heads=4, num_key_value_heads=2, max_position_embeddings=512),
    "llama-160m": dict(vocab_size=32000, hidden_size=768, intermediate_size=2048, num_hidden_layers=12,
                       num_attention_heads=12, num_key_value_heads=4, max_position_embeddings=2048),
}


def llama_config(name: str, **overrides) -> LlamaConfig:
    kw = dict(PRESETS[name])
    kw.update(overrides)
    return LlamaConfig(**kw)


class LlamaDecoderLayer(nn.Module):
    # every matrix parameter of this block is consumed through ops.functional.linear only: ZeRO-3 may hand them to
    # the GEMMs while their all-gather is still in flight (gated B operand, runtime/zero/stage3.py: K2)
    @property
    def _sxb_linear_only_weights(self) -> bool:
        return getattr(self.cfg, "ulysses", None) is None     # the fused Ulysses block reads wqkv / wo outside F.linear

    def __init__(self, cfg: LlamaConfig, layer_idx: int, device=None, dtype=None):
        super().__init__()
        kw = dict(device=device, dtype=dtype)
        h, f, d = cfg.hidden_size, cfg.intermediate_size, cfg.head_dim
        self.cfg = cfg
        self.layer_idx = layer_idx
        self.n_q, self.n_kv, self.d = cfg.num_attention_heads, cfg.num_key_value_heads, d
        self.input_layernorm = nn.Parameter(torch.ones(h, **kw))
        self.wqkv = nn.Parameter(torch.empty((self.n_q + 2 * self.n_kv) * d, h, **kw))
        self.wo = nn.Parameter(torch.empty(h, self.n_q * d, **kw))
        self.post_attention_layernorm = nn.Parameter(torch.ones(h, **kw))
        self.w_gate_up = nn.Parameter(torch.empty(2 * f, h, **kw))
        self.w_down = nn.Parameter(torch.empty(h, f, **kw))

    def reset_parameters(self, std: float):
        for w in (self.wqkv, self.wo, self.w_gate_up, self.w_down):
            if w.device.type != "meta":
                nn.init.normal_(w, mean=0.0, std=std)

    def forward(self, h: torch.Tensor, residual: Optional[torch.Tensor], positions, cos, sin):
        B, S, H = h.shape
        eps = self.cfg.rms_norm_eps
        x, residual = F.rms_norm(h, self.input_layernorm, eps, residual)
        sp = getattr(self.cfg, "ulysses", None)
        if sp is not None:
            # sequence parallel (Ulysses): h holds THIS rank's S tokens of a longer sequence; the attention block
            # exchanges heads <-> sequence inside its GEMM / attention epilogues (sequence/fused_ulysses.py)
            from ..sequence.fused_ulysses import ulysses_attention
            assert B == 1, "the fused Ulysses block takes one sequence per rank"
            h = ulysses_attention(x.view(S, H), self.wqkv, self.wo, cos, sin, sp, True).view(B, S, H)
        else:
            qkv = F.linear(x, self.wqkv)
            qkv = F.rope(qkv, positions, cos, sin, self.n_q + self.n_kv, self.d)
            o = F.flash_attention_qkv(qkv, self.n_q, self.n_kv, self.d, causal=True)
            h = F.linear(o.reshape(B, S, self.n_q * self.d), self.wo)
        x, residual = F.rms_norm(h, self.post_attention_layernorm, eps, residual)
        a = F.swiglu(F.linear(x, self.w_gate_up))
        h = F.linear(a, self.w_down)
        return h, residual


class LlamaModel(nn.Module):
    def __init__(self, cfg: LlamaConfig, device=None, dtype=None):
        super().__init__()
        self.cfg = cfg
        kw = dict(device=device, dtype=dtype)
        self.embed_tokens = nn.Parameter(torch.empty(cfg.vocab_size, cfg.hidden_size, **kw))
        self.layers = nn.ModuleList([LlamaDecoderLayer(cfg, i, **kw) for i in range(cfg.num_hidden_layers)])
        self.norm = nn.Parameter(torch.ones(cfg.hidden_size, **kw))
        self._pos_cache = {}

    def _positions(self, B, S, device):
        key = (B, S, str(device))
        p = self._pos_cache.get(key)
        if p is None:
            p = torch.arange(S, device=device, dtype=torch.long).repeat(B)
            self._pos_cache = {key: p}
        return p

    def forward(self, input_ids: torch.Tensor, position_ids: Optional[torch.Tensor] = None):
        B, S = input_ids.shape
        cfg = self.cfg
        dev = input_ids.device
        positions = position_ids.reshape(-1) if position_ids is not None else self._positions(B, S, dev)
        sp = getattr(cfg, "ulysses", None)
        max_pos = max(cfg.max_position_embeddings, S if sp is None else sp.S)
        cos, sin = F.rope_tables(max_pos, cfg.head_dim, cfg.rope_theta, dev)
        h = F.embedding(input_ids, self.embed_tokens)
        residual = None
        ckpt = cfg.activation_checkpointing and self.training and torch.is_grad_enabled()
        n_ckpt = len(self.layers) if cfg.checkpoint_layers < 0 else cfg.checkpoint_layers
        for i, layer in enumerate(self.layers):
            if ckpt and i < n_ckpt:
                from ..runtime.activation_checkpointing import checkpointing
                h, residual = checkpointing.checkpoint(layer, h, residual, positions, cos, sin)
            else:
                h, residual = layer(h, residual, positions, cos, sin)
        x, _ = F.rms_norm(h, self.norm, cfg.rms_norm_eps, residual)
        return x


class LlamaForCausalLM(nn.Module):
    def __init__(self, cfg: LlamaConfig, device=None, dtype=None):
        super().__init__()
        self.cfg = self.config = cfg
        self.model = LlamaModel(cfg, device=device, dtype=dtype)
        if cfg.tie_word_embeddings:
            self.lm_head = None
        else:
            self.lm_head = nn.Parameter(torch.empty(cfg.vocab_size, cfg.hidden_size, device=device, dtype=dtype))
        self.reset_parameters()

    def reset_parameters(self):
        std = self.cfg.initializer_range
        if self.model.embed_tokens.device.type != "meta":
            nn.init.normal_(self.model.embed_tokens, 0.0, std)
            if self.lm_head is not None:
                nn.init.normal_(self.lm_head, 0.0, std)
        for l in self.model.layers:
            l.reset_parameters(std)

    def head_weight(self):
        return self.lm_head if self.lm_head is not None else self.model.embed_tokens

    def forward(self, input_ids, labels=None, position_ids=None, shift_labels: bool = True):
        x = self.model(input_ids, position_ids)
        logits = F.linear(x, self.head_weight())
        if labels is None:
            return logits
        if shift_labels:
            tgt = torch.full_like(labels, -100)
            tgt[:, :-1] = labels[:, 1:]
        else:
            tgt = labels
        return F.cross_entropy(logits, tgt, ignore_index=-100)

    # --------------------------------------------------------------------------- HF interop
    def load_hf_state_dict(self, sd, strict: bool = True):
        """Accept ``LlamaForCausalLM`` (transformers) parameter names; fuses q/k/v and gate/up."""
        own = {}
        own["model.embed_tokens"] = sd["model.embed_tokens.weight"]
        own["model.norm"] = sd["model.norm.weight"]
        if self.lm_head is not None:
            own["lm_head"] = sd.get("lm_head.weight", sd["model.embed_tokens.weight"])
        for i in range(self.cfg.num_hidden_layers):
            p = f"model.layers.{i}."
            own[p + "input_layernorm"] = sd[p + "input_layernorm.weight"]
            own[p + "post_attention_layernorm"] = sd[p + "post_attention_layernorm.weight"]
            own[p + "wqkv"] = torch.cat([sd[p + "self_attn.q_proj.weight"], sd[p + "self_attn.k_proj.weight"],
                                         sd[p + "self_attn.v_proj.weight"]], 0)
            own[p + "wo"] = sd[p + "self_attn.o_proj.weight"]
            own[p + "w_gate_up"] = torch.cat([sd[p + "mlp.gate_proj.weight"], sd[p + "mlp.up_proj.weight"]], 0)
            own[p + "w_down"] = sd[p + "mlp.down_proj.weight"]
        return self.load_state_dict(own, strict=strict)

    def to_hf_state_dict(self):
        out = {}
        sd = self.state_dict()
        cfg = self.cfg
        d, nq, nk, f = cfg.head_dim, cfg.num_attention_heads, cfg.num_key_value_heads, cfg.intermediate_size
        out["model.embed_tokens.weight"] = sd["model.embed_tokens"]
        out["model.norm.weight"] = sd["model.norm"]
        out["lm_head.weight"] = sd["lm_head"] if "lm_head" in sd else sd["model.embed_tokens"]
        for i in range(cfg.num_hidden_layers):
            p = f"model.layers.{i}."
            out[p + "input_layernorm.weight"] = sd[p + "input_layernorm"]
            out[p + "post_attention_layernorm.weight"] = sd[p + "post_attention_layernorm"]
            q, k, v = sd[p + "wqkv"].split([nq * d, nk * d, nk * d], 0)
            out[p + "self_attn.q_proj.weight"], out[p + "self_attn.k_proj.weight"], out[p + "self_attn.v_proj.weight"] = q, k, v
            out[p + "self_attn.o_proj.weight"] = sd[p + "wo"]
            g, u = sd[p + "w_gate_up"].split([f, f], 0)
            out[p + "mlp.gate_proj.weight"], out[p + "mlp.up_proj.weight"] = g, u
            out[p + "mlp.down_proj.weight"] = sd[p + "w_down"]
        return out


def reference_forward(model: LlamaForCausalLM, input_ids, labels=None):
    """fp32 PyTorch oracle of the same network (used by the numerics tests)."""
    cfg = model.cfg
    sd = {k: v.detach().float() for k, v in model.state_dict().items()}
    B, S = input_ids.shape
    d, nq, nk = cfg.head_dim, cfg.num_attention_heads, cfg.num_key_value_heads
    h = sd["model.embed_tokens"][input_ids]
    inv = 1.0 / (cfg.rope_theta ** (torch.arange(0, d, 2, dtype=torch.float64, device=h.device) / d))
    ang = torch.outer(torch.arange(S, dtype=torch.float64, device=h.device), inv)
    cos, sin = ang.cos().float(), ang.sin().float()

    def rms(x, w):
        return x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + cfg.rms_norm_eps) * w

    def rot(x):  # [B,S,n,d]
        a, b = x[..., : d // 2], x[..., d // 2:]
        c, s = cos[None, :, None, :], sin[None, :, None, :]
        return torch.cat([a * c - b * s, b * c + a * s], -1)

    for i in range(cfg.num_hidden_layers):
        p = f"model.layers.{i}."
        x = rms(h, sd[p + "input_layernorm"])
        qkv = x @ sd[p + "wqkv"].t()
        q = rot(qkv[..., : nq * d].view(B, S, nq, d))
        k = rot(qkv[..., nq * d: (nq + nk) * d].view(B, S, nk, d))
        v = qkv[..., (nq + nk) * d:].view(B, S, nk, d)
        o = F.attention_ref(q, k, v, True)
        h = h + o.reshape(B, S, nq * d) @ sd[p + "wo"].t()
        x = rms(h, sd[p + "post_attention_layernorm"])
        gu = x @ sd[p + "w_gate_up"].t()
        f = cfg.intermediate_size
        h = h + (torch.nn.functional.silu(gu[..., :f]) * gu[..., f:]) @ sd[p + "w_down"].t()
    x = rms(h, sd["model.norm"])
    head = sd["lm_head"] if "lm_head" in sd else sd["model.embed_tokens"]
    logits = x @ head.t()
    if labels is None:
        return logits
    tgt = torch.full_like(labels, -100)
    tgt[:, :-1] = labels[:, 1:]
    return torch.nn.functional.cross_entropy(logits.view(-1, logits.shape[-1]), tgt.view(-1), ignore_index=-100)
