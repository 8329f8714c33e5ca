"""LLaMA / LLaMA-2 / Mistral-style decoder layers (reference: ``containers/llama.py``, ``llama2.py``): RMSNorm, rotary (rotate-half), GQA, gated SiLU MLP, no biases."""
from __future__ import annotations

from ...utils.types import ActivationFuncType, NormType
from .base import AttentionTensors, BaseTransformerContainer, MLPTensors, NormTensors, TransformerPolicy

class LLAMALayerPolicy(TransformerPolicy):
    _orig_layer_class_name = "LlamaDecoderLayer"
    norm_type = NormType.RMSNorm
    mlp_act_func_type = ActivationFuncType.GATED_SILU
    rotate_half = True

    def _cfg(self):
        return self.client_module.self_attn.config

    def get_hidden_heads(self):
        c, cfg = self.client_module, self._cfg()
        eps = getattr(c.input_layernorm, "variance_epsilon", getattr(c.input_layernorm, "eps", 1e-6))
        return cfg.hidden_size, cfg.num_attention_heads, eps, c.mlp.gate_proj.out_features

    def num_kv_heads(self):
        cfg = self._cfg()
        kv = getattr(cfg, "num_key_value_heads", None) or cfg.num_attention_heads
        return -1 if kv == cfg.num_attention_heads else kv

    def rotary(self):
        cfg = self._cfg()
        d = getattr(cfg, "head_dim", None) or cfg.hidden_size // cfg.num_attention_heads
        theta = getattr(cfg, "rope_theta", None)
        if theta is None:
            theta = (getattr(cfg, "rope_parameters", None) or {}).get("rope_theta", 10000.0)
        return d, float(theta)

    def attention(self):
        a = self.client_module.self_attn
        (qw, qb), (kw, kb), (vw, vb), (ow, ob) = (self._wb(m) for m in (a.q_proj, a.k_proj, a.v_proj, a.o_proj))
        return AttentionTensors(qw=qw, qb=qb, kw=kw, kb=kb, vw=vw, vb=vb, ow=ow, ob=ob)

    def mlp(self):
        m = self.client_module.mlp
        return MLPTensors(up_w=m.up_proj.weight.data, gate_w=m.gate_proj.weight.data, out_w=m.down_proj.weight.data,
                          out_b=self._wb(m.down_proj)[1])

    def layernorm(self):
        c = self.client_module
        return NormTensors(attn_nw=c.post_attention_layernorm.weight.data, input_nw=c.input_layernorm.weight.data)


class LLAMA2LayerPolicy(LLAMALayerPolicy):
    """Meta-format checkpoints (``attention.wq`` …, ``feed_forward.w1/w2/w3``) wrapped in a ``TransformerBlock``-like
    module; HuggingFace-format LLaMA-2 goes through :class:`LLAMALayerPolicy`."""
    _orig_layer_class_name = "LlamaTransformerBlock"

    def get_hidden_heads(self):
        c = self.client_module
        return c.attention.wq.weight.shape[1], c.attention.n_local_heads, c.ffn_norm.eps, c.feed_forward.w1.weight.shape[0]

    def num_kv_heads(self):
        a = self.client_module.attention
        kv = getattr(a, "n_local_kv_heads", a.n_local_heads)
        return -1 if kv == a.n_local_heads else kv

    def rotary(self):
        a = self.client_module.attention
        return a.wq.weight.shape[0] // a.n_local_heads, 10000.0

    def attention(self):
        a = self.client_module.attention
        return AttentionTensors(qw=a.wq.weight.data, kw=a.wk.weight.data, vw=a.wv.weight.data, ow=a.wo.weight.data)

    def mlp(self):
        f = self.client_module.feed_forward
        return MLPTensors(up_w=f.w3.weight.data, gate_w=f.w1.weight.data, out_w=f.w2.weight.data)

    def layernorm(self):
        c = self.client_module
        return NormTensors(attn_nw=c.ffn_norm.weight.data, input_nw=c.attention_norm.weight.data)


class DS_LLAMAContainer(BaseTransformerContainer):
    pass


class DS_LLAMA2Container(BaseTransformerContainer):
    pass
