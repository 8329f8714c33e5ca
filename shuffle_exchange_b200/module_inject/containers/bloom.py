"""BLOOM (reference: ``containers/bloom.py``): per-head interleaved fused QKV, ALiBi position biases, pre-LN."""
from __future__ import annotations

from ...utils.types import ActivationFuncType, NormType
from .base import AttentionTensors, BaseTransformerContainer, MLPTensors, NormTensors, TransformerPolicy

class BLOOMLayerPolicy(TransformerPolicy):
    returns_tuple = True          # the HF model indexes ``outputs[0]``
    _orig_layer_class_name = "BloomBlock"

    def get_hidden_heads(self):
        c = self.client_module
        a = c.self_attention
        return a.hidden_size, a.num_heads, c.input_layernorm.eps, c.mlp.dense_h_to_4h.out_features

    def attention(self):
        a = self.client_module.self_attention
        w, b = self._wb(a.query_key_value)
        w, b = self._deinterleave_qkv(w, b, a.num_heads)
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


class DS_BloomContainer(BaseTransformerContainer):
    pass
