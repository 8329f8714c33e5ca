"""GPT-NeoX (reference: ``containers/gptneox.py``): per-head interleaved fused QKV, partial rotary (rotate-half), parallel residual with two LayerNorms."""
from __future__ import annotations

from ...utils.types import ActivationFuncType, NormType
from .base import AttentionTensors, BaseTransformerContainer, MLPTensors, NormTensors, TransformerPolicy

class GPTNEOXLayerPolicy(TransformerPolicy):
    _orig_layer_class_name = "GPTNeoXLayer"
    rotate_half = True

    def __init__(self, client_module=None, inference=True):
        super().__init__(client_module, inference)
        if client_module is not None:
            self.mlp_after_attn = not bool(getattr(client_module, "use_parallel_residual", True))

    def _cfg(self):
        return self.client_module.attention.config

    def get_hidden_heads(self):
        c, cfg = self.client_module, self._cfg()
        return cfg.hidden_size, cfg.num_attention_heads, c.input_layernorm.eps, c.mlp.dense_h_to_4h.out_features

    def rotary(self):
        cfg = self._cfg()
        d = cfg.hidden_size // cfg.num_attention_heads
        pct = getattr(cfg, "rotary_pct", None)
        rp = getattr(cfg, "rope_parameters", None) or {}
        if pct is None:
            pct = rp.get("partial_rotary_factor", 1.0)
        theta = getattr(cfg, "rotary_emb_base", None) or rp.get("rope_theta", 10000.0)
        return int(d * pct), float(theta)

    def attention(self):
        a = self.client_module.attention
        w, b = self._wb(a.query_key_value)
        w, b = self._deinterleave_qkv(w, b, self._cfg().num_attention_heads)
        ow, ob = self._wb(a.dense)
        return AttentionTensors(qkvw=w, qkvb=b, ow=ow, ob=ob)

    def mlp(self):
        m = self.client_module.mlp
        (iw, ib), (ow, ob) = self._wb(m.dense_h_to_4h), self._wb(m.dense_4h_to_h)
        return MLPTensors(inter_w=iw, inter_b=ib, out_w=ow, out_b=ob)

    def layernorm(self):
        c = self.client_module
        return NormTensors(attn_nw=c.post_attention_layernorm.weight.data, attn_nb=c.post_attention_layernorm.bias.data,
                           input_nw=c.input_layernorm.weight.data, input_nb=c.input_layernorm.bias.data)


class DS_GPTNEOXContainer(BaseTransformerContainer):
    pass
