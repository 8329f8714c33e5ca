"""CLIP text/vision encoder layers (reference: ``containers/clip.py``): pre-LN, causal text mask supplied by the caller."""
from __future__ import annotations

from ...utils.types import ActivationFuncType, NormType
from .base import AttentionTensors, BaseTransformerContainer, MLPTensors, NormTensors, TransformerPolicy

class HFCLIPLayerPolicy(TransformerPolicy):
    _orig_layer_class_name = "CLIPEncoderLayer"
    triangular_masking = False       # the text encoder passes its causal mask explicitly; the vision tower has none

    def get_hidden_heads(self):
        c = self.client_module
        return c.self_attn.embed_dim, c.self_attn.num_heads, c.layer_norm1.eps, c.mlp.fc1.out_features

    def attention(self):
        a = self.client_module.self_attn
        (qw, qb), (kw, kb), (vw, vb), (ow, ob) = (self._wb(m) for m in (a.q_proj, a.k_proj, a.v_proj, a.out_proj))
        return AttentionTensors(qw=qw, qb=qb, kw=kw, kb=kb, vw=vw, vb=vb, ow=ow, ob=ob)

    def mlp(self):
        m = self.client_module.mlp
        (iw, ib), (ow, ob) = self._wb(m.fc1), self._wb(m.fc2)
        return MLPTensors(inter_w=iw, inter_b=ib, out_w=ow, out_b=ob)

    def layernorm(self):
        c = self.client_module
        return NormTensors(attn_nw=c.layer_norm2.weight.data, attn_nb=c.layer_norm2.bias.data,
                           input_nw=c.layer_norm1.weight.data, input_nb=c.layer_norm1.bias.data)


class DS_CLIPContainer(BaseTransformerContainer):
    pass
