"""BERT / RoBERTa encoder layers (reference: ``containers/bert.py``): post-LN, bidirectional, split q/k/v."""
from __future__ import annotations

from ...utils.types import ActivationFuncType, NormType
from .base import AttentionTensors, BaseTransformerContainer, MLPTensors, NormTensors, TransformerPolicy

class HFBertLayerPolicy(TransformerPolicy):
    _orig_layer_class_name = "BertLayer"
    pre_attn_norm = False
    triangular_masking = False

    @classmethod
    def match(cls, module):
        return type(module).__name__ in ("BertLayer", "RobertaLayer")

    def get_hidden_heads(self):
        c = self.client_module
        s = c.attention.self
        return s.query.weight.shape[1], s.num_attention_heads, c.attention.output.LayerNorm.eps, \
            c.intermediate.dense.out_features

    def attention(self):
        c = self.client_module
        s = c.attention.self
        (qw, qb), (kw, kb), (vw, vb) = (self._wb(m) for m in (s.query, s.key, s.value))
        ow, ob = self._wb(c.attention.output.dense)
        return AttentionTensors(qw=qw, qb=qb, kw=kw, kb=kb, vw=vw, vb=vb, ow=ow, ob=ob)

    def mlp(self):
        c = self.client_module
        (iw, ib), (ow, ob) = self._wb(c.intermediate.dense), self._wb(c.output.dense)
        return MLPTensors(inter_w=iw, inter_b=ib, out_w=ow, out_b=ob)

    def layernorm(self):
        c = self.client_module
        a, o = c.attention.output.LayerNorm, c.output.LayerNorm
        return NormTensors(attn_nw=a.weight.data, attn_nb=a.bias.data, input_nw=o.weight.data, input_nb=o.bias.data)


class DS_BERTContainer(BaseTransformerContainer):
    pass
