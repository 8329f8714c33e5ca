"""Op objects of the v1 inference transformer.

The reference binds ~20 native entry points per dtype through one class per op (``ops/transformer/inference/
op_binding/*.py``: ``QKVGemmOp``, ``MLPGemmOp``, ``SoftmaxContextOp`` …), each with a PyTorch fallback.  The same op
classes exist here as thin callables over the framework's kernels — tcgen05 GEMM (``F.linear``), fused
LayerNorm/RMSNorm (+residual), fused bias-activation, gated activation, flash attention — so the layer modules
(``ds_attention``, ``ds_mlp``, ``moe_inference``) read like the reference's.  Weights follow the reference's v1 layout:
``[in, out]`` unless ``config.transposed_mode``."""
from __future__ import annotations

import math
from typing import Optional

import torch

from .... import functional as F
from ....fused_softmax import biased_softmax
from .....utils.types import ActivationFuncType, NormType


def dequantize_int8(weight: torch.Tensor, dtype: torch.dtype) -> torch.Tensor:
    """An int8 weight written by ``module_inject.replace_module.GroupQuantizer`` (symmetric, ``weight.scale`` holds one
    step size per group of consecutive rows) back in ``dtype``."""
    scale = weight.scale.to(device=weight.device, dtype=torch.float32).reshape(-1)
    w = weight.float().reshape(scale.numel(), -1) * scale[:, None]
    return w.reshape(weight.shape).to(dtype)


def _w(weight: torch.Tensor, transposed_mode: bool, like: torch.Tensor = None) -> torch.Tensor:
    """v1 stores linear weights as [in, out]; ``F.linear`` wants [out, in].  int8 weights (``dtype=torch.int8``
    injection) are expanded to the activation dtype here, per call: the resident copy stays one byte per element."""
    if weight.dtype == torch.int8:
        weight = dequantize_int8(weight, like.dtype if like is not None else torch.float16)
    return weight if transposed_mode else weight.t()


class BaseOp(torch.nn.Module):

    def __init__(self, config=None):
        super().__init__()
        self.config = config

    @property
    def tm(self) -> bool:
        return bool(getattr(self.config, "transposed_mode", False))


class LayerNormOp(BaseOp):
    def forward(self, vals, gamma, beta, epsilon):
        return F.layer_norm(vals, gamma, beta, epsilon)[0]

    @staticmethod
    def layer_norm_residual(vals, bias, res, gamma, beta, epsilon):
        x = vals if bias is None else vals + bias
        return F.layer_norm(x, gamma, beta, epsilon, residual=res)[0]

    @staticmethod
    def layer_norm_residual_store_pre_ln_res(vals, bias, res, gamma, beta, epsilon):
        x = vals if bias is None else vals + bias
        y, pre = F.layer_norm(x, gamma, beta, epsilon, residual=res)
        return y, pre


class RMSNormOp(BaseOp):
    def forward(self, vals, gamma, epsilon):
        return F.rms_norm(vals, gamma, epsilon)[0]


class PreRMSNormOp(BaseOp):
    def forward(self, vals, residual, gamma, epsilon):
        return F.rms_norm(vals, gamma, epsilon, residual=residual)


class LinearOp(BaseOp):
    def forward(self, input, weight, bias=None, add_bias=True, do_flash_attn=False, num_heads=1, external_cache=None,
                num_layers=None):
        return F.linear(input, _w(weight, self.tm, input), bias if add_bias else None)


class VectorMatMulOp(BaseOp):
    def forward(self, input, weight, async_op=False):
        return F.linear(input, _w(weight, self.tm, input))


class QKVGemmOp(BaseOp):
    """norm(input) @ Wqkv (+ bias); returns ``(qkv, normed_input)``."""

    def forward(self, input, weight, bias, gamma, beta):
        c = self.config
        if c.norm_type == NormType.RMSNorm:
            n = F.rms_norm(input, gamma, c.epsilon)[0]
        elif c.pre_layer_norm:
            n = F.layer_norm(input, gamma, beta, c.epsilon)[0]
        else:
            n = input
        return F.linear(n, _w(weight, self.tm, n), bias), n


class GatedActivationOp(BaseOp):
    def forward(self, activation, bias, activation_func_type):
        x = activation if bias is None else activation + bias
        # reference layout: [.., 2*I] with (up | gate) halves -> act(gate) * up
        return F.gated_act(torch.cat(x.chunk(2, dim=-1)[::-1], dim=-1),
                           "gelu" if activation_func_type == ActivationFuncType.GATED_GELU else "silu")


def _activate(x, bias, act):
    if act in (ActivationFuncType.GATED_GELU, ActivationFuncType.GATED_SILU):
        return GatedActivationOp()(x, bias, act)
    if act == ActivationFuncType.ReLU:
        return F.bias_act(x, bias, "relu")
    return F.bias_act(x, bias, "gelu")


class MLPGemmOp(BaseOp):
    """residual + attn_out (+ attn bias) -> norm -> FFN-in -> activation -> FFN-out; returns ``(out, residual_add)``
    where ``residual_add`` is the pre-norm sum the final residual op needs."""

    def forward(self, input, residual, weight_interm, weight_out, input_bias, bias, gamma, beta):
        c = self.config
        x = input if input_bias is None else input + input_bias
        if not c.mlp_after_attn:
            # parallel residual with its own MLP norm (GPT-NeoX): the MLP reads norm(layer input), not the attention sum
            x, residual = residual, None
        if c.norm_type == NormType.RMSNorm:
            normed, pre = F.rms_norm(x, gamma, c.epsilon, residual=residual)
        else:
            normed, pre = F.layer_norm(x, gamma, beta, c.epsilon, residual=residual)
        inter = _activate(F.linear(normed, _w(weight_interm, self.tm, normed)), bias, c.mlp_act_func_type)
        out = F.linear(inter, _w(weight_out, self.tm, inter))
        if not c.mlp_after_attn:
            return out, None
        # the value the final residual adds to: the pre-norm sum for pre-LN blocks, the normed sum for post-LN blocks
        return out, (pre if c.pre_layer_norm else normed)


class GELUGemmOp(BaseOp):
    def forward(self, input, weight, bias, weight_out):
        inter = F.bias_act(F.linear(input, _w(weight, self.tm, input)), bias, "gelu")
        return F.linear(inter, _w(weight_out, self.tm, inter))


class ResidualAddOp(BaseOp):
    def forward(self, hidden_state, residual, add_bias, attention_output=None, attention_bias=None, final_bias=None,
                residual_add=None):
        c = self.config
        if c.mlp_after_attn:
            base = residual_add if residual_add is not None else residual
            if residual_add is None and attention_output is not None:
                base = residual + attention_output + (attention_bias if attention_bias is not None else 0)
            out = hidden_state + base
            if final_bias is not None:
                out = out + final_bias
            return out
        # parallel attention + MLP (GPT-J / NeoX): everything hangs off the same residual
        out = hidden_state + residual
        if attention_output is not None:
            out = out + attention_output
        if add_bias and attention_bias is not None:
            out = out + attention_bias
        if final_bias is not None:
            out = out + final_bias
        return out


class BiasAddOp(BaseOp):
    def forward(self, activation, bias):
        return activation + bias


class BiasGeluOp(BaseOp):
    def forward(self, activation, bias):
        return F.bias_act(activation, bias, "gelu")


class BiasReluOp(BaseOp):
    def forward(self, activation, bias):
        return F.bias_act(activation, bias, "relu")


class BiasResidualOp(BaseOp):
    def forward(self, output, residual, bias):
        return output + residual + bias


class VectorAddOp(BaseOp):
    def forward(self, a, b, gamma):
        return a * gamma + b


class SoftmaxOp(BaseOp):
    def forward(self, attn_scores, attn_mask, alibi, triangular, recompute, local_attention, window_size,
                async_op, layer_scale, head_offset):
        s = attn_scores * layer_scale
        if alibi is not None:
            s = s + alibi.view(s.shape[0], -1, 1, s.shape[-1])[:, head_offset:head_offset + s.shape[1]]
        if attn_mask is not None:
            s = s + attn_mask
        if triangular:
            q, k = s.shape[-2], s.shape[-1]
            tri = torch.ones(q, k, dtype=torch.bool, device=s.device).tril(k - q)
            s = s.masked_fill(~tri, torch.finfo(s.dtype).min)
        if local_attention:
            q, k = s.shape[-2], s.shape[-1]
            idx = torch.arange(k, device=s.device)[None, :] - (torch.arange(q, device=s.device)[:, None] + k - q)
            s = s.masked_fill(idx.abs() > window_size, torch.finfo(s.dtype).min)
        return torch.softmax(s.float(), dim=-1).to(attn_scores.dtype)


class SoftmaxContextOp(BaseOp):
    """Attention over (past + new) keys for a fused ``qkv [B, S, (H + 2 Hkv) * D]``: applies rotary when configured,
    appends to the layer's KV cache and returns ``(context [B, S, H*D], key, value)``."""

    def forward(self, query_key_value, attn_mask, heads, num_kv, norm_factor, no_masking, layer_id, num_layers,
                alibi, is_prompt=True, token_idx=None, position_ids=None, layer_past=None):
        c = self.config
        B, S, W = query_key_value.shape
        Hkv = heads if num_kv is None or num_kv <= 0 else num_kv
        D = W // (heads + 2 * Hkv)
        q, k, v = query_key_value.split([heads * D, Hkv * D, Hkv * D], dim=-1)
        q, k, v = q.reshape(B, S, heads, D), k.reshape(B, S, Hkv, D), v.reshape(B, S, Hkv, D)
        past = 0 if layer_past is None else layer_past[0].shape[1]
        if getattr(c, "rotary_dim", -1) > 0:
            q, k = _rotary(q, k, c.rotary_dim, past, c.rope_theta, c.rotate_half)
        if layer_past is not None:
            k = torch.cat([layer_past[0], k], dim=1)
            v = torch.cat([layer_past[1], v], dim=1)
        scale = (1.0 / norm_factor) ** 2 if norm_factor else 1.0 / math.sqrt(D)
        causal = c.triangular_masking and not no_masking
        plain = attn_mask is None and alibi is None and past == 0 and q.is_cuda and q.dtype == torch.bfloat16 \
            and D in (64, 128)
        # additive key-padding mask and / or ALiBi (one bias per head and key): handled inside the flash kernels
        key_mask = None
        mask_ok = attn_mask is None or (attn_mask.numel() == B * S and attn_mask.shape[0] == B and attn_mask.shape[-1] == S)
        if past == 0 and mask_ok and (attn_mask is not None or alibi is not None) and q.is_cuda and S % 4 == 0:
            kb = None if attn_mask is None else attn_mask.reshape(B, 1, S).float()
            if alibi is not None:
                al = alibi.reshape(B, heads, -1)[..., :S].float()
                kb = al if kb is None else kb + al
            kb = kb.reshape(B, S) if kb.shape[1] == 1 else kb.contiguous()
            if F.key_bias_supported(q, kb):
                key_mask = kb
        if plain:
            ctx = F.flash_attention(q, k, v, causal=causal, scale=scale)
        elif key_mask is not None:
            ctx = F.flash_attention_key_bias(q, k, v, key_mask, causal=causal, scale=scale)
        else:
            qh = q.permute(0, 2, 1, 3)
            kh = k.permute(0, 2, 1, 3).repeat_interleave(heads // Hkv, dim=1)
            vh = v.permute(0, 2, 1, 3).repeat_interleave(heads // Hkv, dim=1)
            scores = torch.matmul(qh, kh.transpose(-1, -2))
            bias = None
            if causal and S > 1:
                T = k.shape[1]
                tri = torch.ones(S, T, dtype=torch.bool, device=q.device).tril(T - S)
                bias = torch.zeros(1, 1, S, T, dtype=q.dtype, device=q.device).masked_fill(~tri, float("-inf"))
            if alibi is not None:
                al = alibi.reshape(B, heads, 1, -1)[..., :k.shape[1]].to(q.dtype)
                bias = al if bias is None else bias + al
            probs = biased_softmax(scores, attn_mask, bias, scale)
            ctx = torch.matmul(probs, vh).permute(0, 2, 1, 3)
        return ctx.reshape(B, S, heads * D), k, v


def _rotary(q, k, rotary_dim, offset, theta, rotate_half):
    S, D = q.shape[1], q.shape[-1]
    pos = torch.arange(offset, offset + S, device=q.device, dtype=torch.float32)
    inv = 1.0 / (theta ** (torch.arange(0, rotary_dim, 2, device=q.device, dtype=torch.float32) / rotary_dim))
    ang = pos[:, None] * inv[None, :]
    cos, sin = ang.cos()[None, :, None, :], ang.sin()[None, :, None, :]

    def rot(x):
        xr, xp = x[..., :rotary_dim].float(), x[..., rotary_dim:]
        if rotate_half:
            x1, x2 = xr[..., :rotary_dim // 2], xr[..., rotary_dim // 2:]
            out = torch.cat([x1 * cos - x2 * sin, x2 * cos + x1 * sin], dim=-1)
        else:      # rotate every two (GPT-J style interleaved pairs)
            x1, x2 = xr[..., 0::2], xr[..., 1::2]
            out = torch.stack([x1 * cos - x2 * sin, x2 * cos + x1 * sin], dim=-1).flatten(-2)
        return torch.cat([out.to(x.dtype), xp], dim=-1)

    return rot(q), rot(k)


class EinsumSecSmEcmOp(BaseOp):
    def forward(self, Q, W):
        return torch.einsum("sec,sm->ecm", Q, W)


class MoEResMatmulOp(BaseOp):
    def forward(self, residual, coef, output):
        c1, c2 = coef.chunk(2, dim=-1)
        return residual * c1 + output * c2


class PadTransformOp(BaseOp):
    def forward(self, query, key, value, heads, add_padding):
        B, S, W = query.shape
        D = W // heads
        f = lambda t: t.reshape(B, S, heads, D).permute(0, 2, 1, 3).contiguous()   # noqa: E731
        return f(query), f(key), f(value)


class WorkspaceOp(BaseOp):
    """KV-cache / scratch arena.  Caches are ordinary tensors owned by the attention modules here, so the workspace is
    bookkeeping only: it records the reservation the reference would make and lets the hybrid engine release it."""

    def __init__(self, config=None):
        super().__init__(config)
        self.reserved_bytes = 0
        self.allocated = False

    def allocate_workspace(self, hidden_size, heads, prompt_len, batch, num_layers, mp_size, external_cache, rank,
                           max_out_tokens, min_out_tokens):
        per_tok = 2 * hidden_size // max(mp_size, 1) * 2
        self.reserved_bytes = batch * (prompt_len + max_out_tokens) * per_tok * num_layers
        self.allocated = True

    def release_workspace(self):
        self.allocated = False
        return True

    def retake_workspace(self):
        self.allocated = True
        return True

    def is_allocated(self):
        return self.allocated

    def reset_cache(self):
        return None
