"""OPT (reference: ``containers/opt.py``): pre-LN (``do_layer_norm_before``), split q/k/v, ReLU (or GELU) MLP."""
from __future__ import annotations

from ...utils.types import ActivationFuncType, NormType
from .base import AttentionTensors, BaseTransformerContainer, MLPTensors, NormTensors, TransformerPolicy

class HFOPTLayerPolicy(TransformerPolicy):
    _orig_layer_class_name = "OPTDecoderLayer"

    def __init__(self, client_module=None, inference=True):
        super().__init__(client_module, inference)
        if client_module is not None:
            self.pre_attn_norm = bool(getattr(client_module, "do_layer_norm_before", True))
            act = type(getattr(client_module, "activation_fn", None)).__name__.lower()
            self.mlp_act_func_type = ActivationFuncType.ReLU if "relu" in act else ActivationFuncType.GELU

    def get_hidden_heads(self):
        c = self.client_module
        return c.self_attn.embed_dim, c.self_attn.num_heads, c.self_attn_layer_norm.eps, c.fc1.out_features

    def attention(self):
        a = self.client_module.self_attn
        (qw, qb), (kw, kb), (vw, vb), (ow, ob) = (self._wb(m) for m in (a.q_proj, a.k_proj, a.v_proj, a.out_proj))
        return AttentionTensors(qw=qw, qb=qb, kw=kw, kb=kb, vw=vw, vb=vb, ow=ow, ob=ob)

    def mlp(self):
        c = self.client_module
        (iw, ib), (ow, ob) = self._wb(c.fc1), self._wb(c.fc2)
        return MLPTensors(inter_w=iw, inter_b=ib, out_w=ow, out_b=ob)

    def layernorm(self):
        c = self.client_module
        return NormTensors(attn_nw=c.final_layer_norm.weight.data, attn_nb=c.final_layer_norm.bias.data,
                           input_nw=c.self_attn_layer_norm.weight.data, input_nb=c.self_attn_layer_norm.bias.data)


class DS_OPTContainer(BaseTransformerContainer):
    pass
