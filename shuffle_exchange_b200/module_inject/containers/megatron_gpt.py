"""Megatron-LM GPT ``ParallelTransformerLayer`` (reference: ``containers/megatron_gpt.py`` + ``features/megatron.py``): fused QKV that is per-head interleaved from checkpoint version 2 on."""
from __future__ import annotations

from ...utils.types import ActivationFuncType, NormType
from .base import AttentionTensors, BaseTransformerContainer, MLPTensors, NormTensors, TransformerPolicy

class MegatronLayerPolicy(TransformerPolicy):
    _orig_layer_class_name = "ParallelTransformerLayer"
    version = 0
    moe_type = "standard"
    megatron_v2 = True
    use_mup = False

    def _attn(self):
        c = self.client_module
        return c.attention if hasattr(c, "attention") else c.self_attention

    def get_hidden_heads(self):
        c, a = self.client_module, self._attn()
        return a.query_key_value.weight.shape[1], a.num_attention_heads, c.input_layernorm.eps, \
            c.mlp.dense_h_to_4h.weight.shape[0]

    def attention(self):
        a = self._attn()
        w, b = self._wb(a.query_key_value)
        if self.megatron_v2:
            w, b = self._deinterleave_qkv(w, b, a.num_attention_heads)
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


class DS_MegatronGPTContainer(BaseTransformerContainer):
    pass
