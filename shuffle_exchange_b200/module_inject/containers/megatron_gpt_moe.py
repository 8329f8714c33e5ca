"""Megatron GPT with MoE MLPs (reference: ``containers/megatron_gpt_moe.py`` + ``base_moe.py``): the attention half is
injected exactly like dense Megatron; expert weights stay in the client's ``MoE`` module, which already runs on the
framework's dropless / capacity MoE kernels."""
from __future__ import annotations

from .base import BaseTransformerContainer, MLPTensors
from .megatron_gpt import MegatronLayerPolicy


class MegatronMoELayerPolicy(MegatronLayerPolicy):
    _orig_layer_class_name = "ParallelTransformerLayerMoE"
    moe_type = "standard"
    num_experts = 1

    def __init__(self, client_module=None, inference=True):
        super().__init__(client_module, inference)
        moe = getattr(getattr(client_module, "mlp", None), "deepspeed_moe", None)
        if moe is not None:
            self.num_experts = getattr(moe, "num_experts", len(getattr(moe.experts, "deepspeed_experts", [])) or 1)

    def get_hidden_heads(self):
        c, a = self.client_module, self._attn()
        experts = c.mlp.deepspeed_moe.experts.deepspeed_experts
        return a.query_key_value.weight.shape[1], a.num_attention_heads, c.input_layernorm.eps, \
            experts[0].dense_h_to_4h.weight.shape[0]

    def mlp(self, moe_type="standard"):
        experts = self.client_module.mlp.deepspeed_moe.experts.deepspeed_experts
        e0 = experts[0]
        return MLPTensors(inter_w=e0.dense_h_to_4h.weight.data, inter_b=e0.dense_h_to_4h.bias.data,
                          out_w=e0.dense_4h_to_h.weight.data, out_b=e0.dense_4h_to_h.bias.data)

    def expert_tensors(self):
        """Per-expert (inter_w, inter_b, out_w, out_b) for ``DeepSpeedMoEInference``."""
        return [(e.dense_h_to_4h.weight.data, e.dense_h_to_4h.bias.data, e.dense_4h_to_h.weight.data,
                 e.dense_4h_to_h.bias.data) for e in self.client_module.mlp.deepspeed_moe.experts.deepspeed_experts]


class BaseTransformerMoEContainer(BaseTransformerContainer):
    """MoE variant (reference ``base_moe.py``): builds ``DeepSpeedMoEInference`` and fills every local expert."""

    def create_module(self, config=None):
        from ...ops.transformer.inference import DeepSpeedMoEInference, DeepSpeedMoEInferenceConfig
        p = self.policy
        cfg = DeepSpeedMoEInferenceConfig(
            hidden_size=self.hidden_size, heads=self.num_attention_heads, layer_norm_eps=self.layernorm_epsilon,
            dtype=self.dtype, pre_layer_norm=p.pre_attn_norm, mp_size=self.mp_size, scale_attention=p.scale_attention,
            moe_experts=p.num_experts, global_experts=p.num_experts, mlp_type=p.moe_type, return_tuple=False)
        self.ds_model_config = cfg
        self.module = DeepSpeedMoEInference(cfg, mp_group=self.mp_group)
        return self.module


class DS_MegatronGPTMoEContainer(BaseTransformerMoEContainer):
    pass
