"""GPT-2 (reference: ``containers/gpt2.py``): pre-LN, fused Conv1D QKV ``[in, out]``, tanh-GELU."""
from __future__ import annotations

from ...utils.types import ActivationFuncType, NormType
from .base import AttentionTensors, BaseTransformerContainer, MLPTensors, NormTensors, TransformerPolicy

class HFGPT2LayerPolicy(TransformerPolicy):
    _orig_layer_class_name = "GPT2Block"
    linear_layer = False

    def get_hidden_heads(self):
        a = self.client_module.attn
        return a.embed_dim, a.num_heads, self.client_module.ln_1.eps, self.client_module.mlp.c_fc.weight.shape[1]

    def attention(self):
        a = self.client_module.attn
        qkvw, qkvb = self._conv1d(a.c_attn)
        ow, ob = self._conv1d(a.c_proj)
        return AttentionTensors(qkvw=qkvw, qkvb=qkvb, ow=ow, ob=ob)

    def mlp(self):
        m = self.client_module.mlp
        iw, ib = self._conv1d(m.c_fc)
        ow, ob = self._conv1d(m.c_proj)
        return MLPTensors(inter_w=iw, inter_b=ib, out_w=ow, out_b=ob)

    def layernorm(self):
        c = self.client_module
        return NormTensors(attn_nw=c.ln_2.weight.data, attn_nb=c.ln_2.bias.data, input_nw=c.ln_1.weight.data,
                           input_nb=c.ln_1.bias.data)


class DS_GPT2Container(BaseTransformerContainer):
    pass
