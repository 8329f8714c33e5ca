"""GPT-Neo (reference: ``containers/gptneo.py``): pre-LN, unscaled attention scores, alternating global / local (windowed) layers."""
from __future__ import annotations

from ...utils.types import ActivationFuncType, NormType
from .base import AttentionTensors, BaseTransformerContainer, MLPTensors, NormTensors, TransformerPolicy

class HFGPTNEOLayerPolicy(TransformerPolicy):
    returns_tuple = True          # the HF model indexes ``outputs[0]``
    _orig_layer_class_name = "GPTNeoBlock"
    scale_attention = False

    def __init__(self, client_module=None, inference=True):
        super().__init__(client_module, inference)
        if client_module is not None:
            self.local_attention = getattr(client_module.attn, "attention_type", "global") == "local"
            cfg = getattr(client_module.attn.attention, "config", None)
            self.window_size = getattr(cfg, "window_size", 256)

    def get_hidden_heads(self):
        c = self.client_module
        a = c.attn.attention
        return a.embed_dim, a.num_heads, c.ln_1.eps, c.mlp.c_fc.out_features

    def attention(self):
        a = self.client_module.attn.attention
        ow, ob = self._wb(a.out_proj)
        return AttentionTensors(qw=a.q_proj.weight.data, kw=a.k_proj.weight.data, vw=a.v_proj.weight.data, ow=ow, ob=ob)

    def mlp(self):
        m = self.client_module.mlp
        (iw, ib), (ow, ob) = self._wb(m.c_fc), self._wb(m.c_proj)
        return MLPTensors(inter_w=iw, inter_b=ib, out_w=ow, out_b=ob)

    def layernorm(self):
        c = self.client_module
        return NormTensors(attn_nw=c.ln_2.weight.data, attn_nb=c.ln_2.bias.data, input_nw=c.ln_1.weight.data,
                           input_nb=c.ln_1.bias.data)


class DS_GPTNEOContainer(BaseTransformerContainer):
    pass
