"""v1 inference transformer layer (attention + MLP with KV cache), MoE inference layer, diffusers ops and graphed
model wrappers — CPU numerics against plain PyTorch."""
import math

import pytest
import torch

from shuffle_exchange_b200.model_implementations import DSUNet, DSVAE, DSClipEncoder, DeepSpeedTransformerInference
from shuffle_exchange_b200.ops.transformer.inference import (DeepSpeedDiffusersAttention, DeepSpeedInferenceConfig,
                                                             DeepSpeedMoEInference, DeepSpeedMoEInferenceConfig,
                                                             nhwc_bias_add)
from shuffle_exchange_b200.utils.types import ActivationFuncType


def _fill(module, std=0.1):
    torch.manual_seed(0)
    for n, p in module.named_parameters():
        if p.dim() == 1 and ("nw" in n or "norm_w" in n):
            p.data.fill_(1.0).add_(torch.randn_like(p) * 0.01)
        else:
            p.data.normal_(0, std)


def _ref_layer(layer, x, past=None):
    c = layer.config
    a, m = layer.attention, layer.mlp
    H, D = c.heads, c.hidden_size // c.heads
    ln = lambda t, w, b: torch.nn.functional.layer_norm(t, (c.hidden_size,), w, b, c.epsilon)
    h = ln(x, layer.norm_w, layer.norm_b)
    qkv = h @ a.attn_qkvw + a.attn_qkvb
    B, S, _ = x.shape
    q, k, v = qkv.split(c.hidden_size, dim=-1)
    q, k, v = (t.view(B, S, H, D).transpose(1, 2) for t in (q, k, v))
    if past is not None:
        k = torch.cat([past[0], k], 2)
        v = torch.cat([past[1], v], 2)
    T = k.shape[2]
    s = q @ k.transpose(-1, -2) / math.sqrt(D)
    mask = torch.ones(S, T, dtype=torch.bool).tril(T - S)
    s = s.masked_fill(~mask, float("-inf"))
    ctx = (torch.softmax(s, -1) @ v).transpose(1, 2).reshape(B, S, -1)
    att = ctx @ a.attn_ow
    res = x + att + a.attn_ob
    h2 = ln(res, m.attn_nw, m.attn_nb)
    out = torch.nn.functional.gelu(h2 @ m.inter_w + m.inter_b, approximate="tanh") @ m.output_w + m.output_b + res
    return out, (k, v)


def test_transformer_inference_layer_prompt_then_decode_matches_reference():
    DeepSpeedTransformerInference.layer_id = 0
    cfg = DeepSpeedInferenceConfig(hidden_size=32, heads=4, num_hidden_layers=1, layer_norm_eps=1e-5,
                                   dtype=torch.float32, pre_layer_norm=True, return_tuple=True)
    layer = DeepSpeedTransformerInference(cfg)
    _fill(layer)
    x = torch.randn(2, 6, 32)
    out, presents = layer(x, use_cache=True)
    ref, (k, v) = _ref_layer(layer, x)
    assert torch.allclose(out, ref, atol=1e-4), (out - ref).abs().max()
    assert presents[0].shape == (2, 6, 4, 8)
    # one decode step reuses the layer-internal cache
    x1 = torch.randn(2, 1, 32)
    out1, presents1 = layer(x1, use_cache=True)
    ref1, _ = _ref_layer(layer, x1, past=(k, v))
    assert torch.allclose(out1, ref1, atol=1e-4)
    assert presents1[0].shape[1] == 7
    # HF-style keyword + explicit past
    out2 = layer(hidden_states=x1, layer_past=presents, use_cache=False)
    assert torch.allclose(out2[0], ref1, atol=1e-4)
    assert DeepSpeedTransformerInference.workspace.is_allocated()


def test_transformer_inference_gated_rms_gqa_rotary_runs_and_is_causal():
    from shuffle_exchange_b200.utils.types import NormType
    DeepSpeedTransformerInference.layer_id = 0
    cfg = DeepSpeedInferenceConfig(hidden_size=32, intermediate_size=48, heads=4, num_kv=2, num_hidden_layers=1,
                                   dtype=torch.float32, norm_type=NormType.RMSNorm, rotary_dim=8, rotate_half=True,
                                   mlp_act_func_type=ActivationFuncType.GATED_SILU, return_tuple=False)
    layer = DeepSpeedTransformerInference(cfg)
    _fill(layer)
    assert layer.attention.attn_qkvw.shape == (32, (4 + 2 * 2) * 8) and layer.mlp.inter_w.shape == (32, 96)
    x = torch.randn(1, 5, 32)
    full = layer(x)
    part = layer(x[:, :3])
    assert torch.allclose(full[:, :3], part, atol=1e-5)          # causal: a prefix is unaffected by later tokens


def test_moe_inference_layer_matches_dense_top1_routing():
    DeepSpeedMoEInference.layer_id = 0
    cfg = DeepSpeedMoEInferenceConfig(hidden_size=16, intermediate_size=32, heads=2, num_hidden_layers=1,
                                      layer_norm_eps=1e-5, moe_experts=4, global_experts=4, k=1, return_tuple=False)
    layer = DeepSpeedMoEInference(cfg)
    _fill(layer)
    x = torch.randn(2, 5, 16)
    out = layer(x)
    with torch.no_grad():
        att = layer.attention(x, None, None, None, False, None, None, False, layer.norm_w, layer.norm_b)[0]
        res = att + layer.attention.attn_ob + x
        h = torch.nn.functional.layer_norm(res, (16,), layer.attn_nw, layer.attn_nb, 1e-5).reshape(-1, 16)
        probs = torch.softmax(h @ layer.moe_gate_w.t(), -1)
        top = probs.argmax(-1)
        y = torch.zeros_like(h)
        for e in range(4):
            sel = top == e
            ex = layer.mlp[e]
            y[sel] = (torch.nn.functional.gelu(h[sel] @ ex.inter_w + ex.inter_b, approximate="tanh") @ ex.output_w
                      + ex.output_b) * probs[sel, e:e + 1]
        ref = y.view(2, 5, 16) + res
    assert torch.allclose(out, ref, atol=1e-4), (out - ref).abs().max()


def test_diffusers_attention_and_bias_add():
    cfg = DeepSpeedInferenceConfig(hidden_size=32, heads=4, dtype=torch.float32)
    att = DeepSpeedDiffusersAttention(cfg)
    _fill(att)
    x = torch.randn(2, 9, 32)
    out = att(x)
    qkv = x @ att.attn_qkvw + att.attn_qkvb
    q, k, v = (t.view(2, 9, 4, 8).transpose(1, 2) for t in qkv.chunk(3, -1))
    ref = (torch.softmax(q @ k.transpose(-1, -2) / math.sqrt(8), -1) @ v).transpose(1, 2).reshape(2, 9, 32) \
        @ att.attn_ow + att.attn_ob
    assert torch.allclose(out, ref, atol=1e-4)
    ctx = torch.randn(2, 4, 32)
    assert att(x, ctx).shape == (2, 9, 32)
    assert att(torch.randn(2, 32, 3, 3)).shape == (2, 32, 3, 3)
    a, b = torch.randn(2, 4, 3, 3), torch.randn(4)
    assert torch.allclose(nhwc_bias_add(a, b), a + b.view(1, 4, 1, 1))
    assert torch.allclose(nhwc_bias_add(a, b, a, b), 2 * (a + b.view(1, 4, 1, 1)))


def test_graphed_wrappers_fall_back_to_eager_without_cuda():
    class UNet(torch.nn.Module):
        in_channels = 4
        def __init__(self):
            super().__init__()
            self.conv = torch.nn.Conv2d(4, 4, 1)
        def forward(self, sample, t, enc, return_dict=True, **kw):
            return self.conv(sample) + t

    class VAE(torch.nn.Module):
        def decode(self, x, return_dict=True): return x * 2
        def encode(self, x, return_dict=True): return x / 2
        def forward(self, s, t=None, e=None, return_dict=True): return s + 1

    u = DSUNet(UNet(), enable_cuda_graph=True)
    s = torch.randn(1, 4, 8, 8)
    assert torch.allclose(u(s, torch.tensor(1.0), None), u.unet(s, torch.tensor(1.0), None))
    with pytest.raises(NotImplementedError):
        u(s, 1, None, cross_attention_kwargs={"scale": 1})
    v = DSVAE(VAE())
    assert torch.equal(v.decode(s), s * 2) and torch.equal(v.encode(s), s / 2) and torch.equal(v(s), s + 1)
    enc = DSClipEncoder(torch.nn.Linear(4, 4))
    x = torch.randn(2, 4)
    assert torch.allclose(enc(x), enc.enc(x)) and enc.iter == 1
    m = enc._build_causal_attention_mask(1, 3, torch.float32)
    assert m.shape == (1, 1, 3, 3) and m[0, 0, 0, 1] < -1e30 and m[0, 0, 1, 0] == 0
