"""GPT-J (reference: ``containers/gptj.py``): one LayerNorm feeding attention and MLP in parallel, interleaved rotary on the first ``rotary_dim`` dims, bias-free attention projections."""
from __future__ import annotations

from ...utils.types import ActivationFuncType, NormType
from .base import AttentionTensors, BaseTransformerContainer, MLPTensors, NormTensors, TransformerPolicy

class HFGPTJLayerPolicy(TransformerPolicy):
    returns_tuple = True          # the HF model indexes ``outputs[0]``
    _orig_layer_class_name = "GPTJBlock"
    mlp_after_attn = False
    rotate_every_two = True

    def get_hidden_heads(self):
        c = self.client_module
        return c.attn.embed_dim, c.attn.num_attention_heads, c.ln_1.eps, c.mlp.fc_in.out_features

    def rotary(self):
        a = self.client_module.attn
        return int(a.rotary_dim or a.head_dim), 10000.0

    def attention(self):
        a = self.client_module.attn
        return AttentionTensors(qw=a.q_proj.weight.data, kw=a.k_proj.weight.data, vw=a.v_proj.weight.data,
                                ow=a.out_proj.weight.data)

    def mlp(self):
        m = self.client_module.mlp
        (iw, ib), (ow, ob) = self._wb(m.fc_in), self._wb(m.fc_out)
        return MLPTensors(inter_w=iw, inter_b=ib, out_w=ow, out_b=ob)

    def layernorm(self):
        c = self.client_module
        return NormTensors(input_nw=c.ln_1.weight.data, input_nb=c.ln_1.bias.data)


class DS_GPTJContainer(BaseTransformerContainer):
    pass
