"""DistilBERT (reference: ``containers/distil_bert.py``): post-LN, bidirectional."""
from __future__ import annotations

from ...utils.types import ActivationFuncType, NormType
from .base import AttentionTensors, BaseTransformerContainer, MLPTensors, NormTensors, TransformerPolicy

class HFDistilBertLayerPolicy(TransformerPolicy):
    _orig_layer_class_name = "TransformerBlock"
    pre_attn_norm = False
    triangular_masking = False

    def get_hidden_heads(self):
        c = self.client_module
        return c.attention.q_lin.weight.shape[1], c.attention.n_heads, c.sa_layer_norm.eps, c.ffn.lin1.out_features

    def attention(self):
        a = self.client_module.attention
        (qw, qb), (kw, kb), (vw, vb), (ow, ob) = (self._wb(m) for m in (a.q_lin, a.k_lin, a.v_lin, a.out_lin))
        return AttentionTensors(qw=qw, qb=qb, kw=kw, kb=kb, vw=vw, vb=vb, ow=ow, ob=ob)

    def mlp(self):
        f = self.client_module.ffn
        (iw, ib), (ow, ob) = self._wb(f.lin1), self._wb(f.lin2)
        return MLPTensors(inter_w=iw, inter_b=ib, out_w=ow, out_b=ob)

    def layernorm(self):
        c = self.client_module
        return NormTensors(attn_nw=c.sa_layer_norm.weight.data, attn_nb=c.sa_layer_norm.bias.data,
                           input_nw=c.output_layer_norm.weight.data, input_nb=c.output_layer_norm.bias.data)


class DS_DistilBERTContainer(BaseTransformerContainer):
    pass
