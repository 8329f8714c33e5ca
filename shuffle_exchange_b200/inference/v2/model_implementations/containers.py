"""Architecture containers written in the declarative DSL (reference: ``<arch>/container.py`` of every v2 model).

They are an alternative, equivalent front end to the rule tables in ``policies.py``: ``ContainerPolicyMixin`` (below)
makes a policy build its weights by streaming checkpoint tensors into these containers.  New architectures can be added
either way; the containers are the reference's extension interface."""
from __future__ import annotations

import re
from typing import Any, Dict, List, Optional

import torch

from .common_parameters import *  # noqa: F401,F403
from .common_parameters import (AttentionOutputParameter, EmbeddingParameter, GatedMLPParameter,
                                GQAMegatronQKVParameter, MLP1Parameter, MLP2Parameter, MoEGatingWeightParameter,
                                NormParameter, UnembedParameter, UnfusedMoEGatedMLPParameter, UnfusedMoEMLP2Parameter,
                                UnfusedQKVParameter)
from .layer_container_base import LayerContainer


# ------------------------------------------------------------------------------------------------ llama family
class Llama2TransformerContainer(LayerContainer):
    qkv_w: UnfusedQKVParameter
    qkv_b: Optional[UnfusedQKVParameter]            # Qwen2 carries q/k/v biases, Llama / Mistral do not
    attn_out_w: AttentionOutputParameter
    mlp_1_w: GatedMLPParameter
    mlp_2_w: MLP2Parameter
    attn_norm_gamma: NormParameter
    mlp_norm_gamma: NormParameter

    PARAM_MAPPING = {
        "self_attn.q_proj.weight": "qkv_w.q_params",
        "self_attn.k_proj.weight": "qkv_w.k_params",
        "self_attn.v_proj.weight": "qkv_w.v_params",
        "self_attn.q_proj.bias": "qkv_b.q_params",
        "self_attn.k_proj.bias": "qkv_b.k_params",
        "self_attn.v_proj.bias": "qkv_b.v_params",
        "self_attn.o_proj.weight": "attn_out_w.params",
        "mlp.gate_proj.weight": "mlp_1_w.gate_params",
        "mlp.up_proj.weight": "mlp_1_w.up_params",
        "mlp.down_proj.weight": "mlp_2_w.params",
        "input_layernorm.weight": "attn_norm_gamma.params",
        "post_attention_layernorm.weight": "mlp_norm_gamma.params",
    }


class Llama2NonTransformerContainer(LayerContainer):
    word_emb: EmbeddingParameter
    word_unembed: Optional[UnembedParameter]        # absent when the embeddings are tied
    final_norm: NormParameter

    PARAM_MAPPING = {
        "model.embed_tokens.weight": "word_emb.params",
        "model.norm.weight": "final_norm.params",
        "lm_head.weight": "word_unembed.params",
    }


MistralTransformerContainer, MistralNonTransformerContainer = Llama2TransformerContainer, Llama2NonTransformerContainer
Qwen2TransformerContainer, Qwen2NonTransformerContainer = Llama2TransformerContainer, Llama2NonTransformerContainer


# ------------------------------------------------------------------------------------------------ mixtral
class MixtralTransformerContainer(LayerContainer):
    qkv_w: UnfusedQKVParameter
    attn_out_w: AttentionOutputParameter
    moe_gate: MoEGatingWeightParameter
    moe_mlp_1: UnfusedMoEGatedMLPParameter
    moe_mlp_2: UnfusedMoEMLP2Parameter
    attn_norm_gamma: NormParameter
    mlp_norm_gamma: NormParameter

    PARAM_MAPPING = {
        "self_attn.q_proj.weight": "qkv_w.q_params",
        "self_attn.k_proj.weight": "qkv_w.k_params",
        "self_attn.v_proj.weight": "qkv_w.v_params",
        "self_attn.o_proj.weight": "attn_out_w.params",
        "block_sparse_moe.gate.weight": "moe_gate.params",
        "block_sparse_moe.experts.*.w1.weight": "moe_mlp_1.gating_experts",
        "block_sparse_moe.experts.*.w3.weight": "moe_mlp_1.up_experts",
        "block_sparse_moe.experts.*.w2.weight": "moe_mlp_2.experts",
        "input_layernorm.weight": "attn_norm_gamma.params",
        "post_attention_layernorm.weight": "mlp_norm_gamma.params",
    }


MixtralNonTransformerContainer = Llama2NonTransformerContainer


# ------------------------------------------------------------------------------------------------ OPT
class OPTTransformerContainer(LayerContainer):
    qkv_w: UnfusedQKVParameter
    qkv_b: UnfusedQKVParameter
    attn_out_w: AttentionOutputParameter
    attn_out_b: AttentionOutputParameter
    mlp_1_w: MLP1Parameter
    mlp_1_b: MLP1Parameter
    mlp_2_w: MLP2Parameter
    mlp_2_b: MLP2Parameter
    attn_norm_gamma: NormParameter
    attn_norm_beta: NormParameter
    mlp_norm_gamma: NormParameter
    mlp_norm_beta: NormParameter

    PARAM_MAPPING = {
        "self_attn.q_proj.weight": "qkv_w.q_params", "self_attn.q_proj.bias": "qkv_b.q_params",
        "self_attn.k_proj.weight": "qkv_w.k_params", "self_attn.k_proj.bias": "qkv_b.k_params",
        "self_attn.v_proj.weight": "qkv_w.v_params", "self_attn.v_proj.bias": "qkv_b.v_params",
        "self_attn.out_proj.weight": "attn_out_w.params", "self_attn.out_proj.bias": "attn_out_b.params",
        "fc1.weight": "mlp_1_w.params", "fc1.bias": "mlp_1_b.params",
        "fc2.weight": "mlp_2_w.params", "fc2.bias": "mlp_2_b.params",
        "self_attn_layer_norm.weight": "attn_norm_gamma.params", "self_attn_layer_norm.bias": "attn_norm_beta.params",
        "final_layer_norm.weight": "mlp_norm_gamma.params", "final_layer_norm.bias": "mlp_norm_beta.params",
    }


class OPTNonTransformerContainer(LayerContainer):
    word_emb: EmbeddingParameter
    word_emb_pos: EmbeddingParameter
    word_unembed: Optional[UnembedParameter]
    final_norm_w: NormParameter
    final_norm_b: NormParameter

    PARAM_MAPPING = {
        "model.decoder.embed_tokens.weight": "word_emb.params",
        "model.decoder.embed_positions.weight": "word_emb_pos.params",
        "model.decoder.final_layer_norm.weight": "final_norm_w.params",
        "model.decoder.final_layer_norm.bias": "final_norm_b.params",
        "lm_head.weight": "word_unembed.params",
    }


# ------------------------------------------------------------------------------------------------ Falcon (new decoder)
class FalconNewArchTransformerContainer(LayerContainer):
    qkv_w: GQAMegatronQKVParameter
    attn_out_w: AttentionOutputParameter
    mlp_1_w: MLP1Parameter
    mlp_2_w: MLP2Parameter
    attn_norm_gamma: NormParameter
    attn_norm_beta: NormParameter
    mlp_norm_gamma: NormParameter
    mlp_norm_beta: NormParameter

    PARAM_MAPPING = {
        "self_attention.query_key_value.weight": "qkv_w.params",
        "self_attention.dense.weight": "attn_out_w.params",
        "mlp.dense_h_to_4h.weight": "mlp_1_w.params",
        "mlp.dense_4h_to_h.weight": "mlp_2_w.params",
        "ln_attn.weight": "attn_norm_gamma.params", "ln_attn.bias": "attn_norm_beta.params",
        "ln_mlp.weight": "mlp_norm_gamma.params", "ln_mlp.bias": "mlp_norm_beta.params",
    }


class FalconNonTransformerContainer(LayerContainer):
    word_emb: EmbeddingParameter
    word_unembed: Optional[UnembedParameter]
    final_norm_w: NormParameter
    final_norm_b: NormParameter

    PARAM_MAPPING = {
        "transformer.word_embeddings.weight": "word_emb.params",
        "transformer.ln_f.weight": "final_norm_w.params",
        "transformer.ln_f.bias": "final_norm_b.params",
        "lm_head.weight": "word_unembed.params",
    }


# ------------------------------------------------------------------------------------------------ policy mix-in
class _ModelView:
    """What parameters may ask of "the model" while assembling: the ArchSpec fields by their model-property names."""

    def __init__(self, spec):
        self.spec = spec
        self.n_heads, self.n_heads_kv, self.head_size = spec.n_heads, spec.n_heads_kv, spec.head_size
        self.n_experts, self.model_dim, self.num_layers = spec.n_experts, spec.model_dim, spec.num_layers


class ContainerPolicyMixin:
    """Put in front of an :class:`InferenceV2Policy` subclass and set ``transformer_container`` /
    ``non_transformer_container``: ``load_weights`` then streams the checkpoint through the containers instead of the
    rule tables.  Unmapped checkpoint names are reported through ``container_map`` exactly as on the rule path."""

    transformer_container: type = None
    non_transformer_container: type = None

    def load_weights(self) -> Dict:
        if self._weights is not None:
            return self._weights
        if self._inf_checkpoint_path is not None:
            return super().load_weights()
        spec = self.resolved_spec()
        view = _ModelView(spec)
        dev, dt = self._device(), spec.dtype
        layers: List[LayerContainer] = [self.transformer_container(view) for _ in range(spec.num_layers)]
        glob = self.non_transformer_container(view)
        pat = re.compile("^" + re.escape(self.layer_prefix) + r"(\d+)\.(.+)$")
        ignore = [re.compile(p) for p in self.ignore_patterns]
        unmapped = []
        for name, t in self._checkpoint_engine.parameters():
            if any(p.search(name) for p in ignore):
                continue
            t = t.detach()
            t = t.to(dev, dt) if t.is_floating_point() else t.to(dev)
            m = pat.match(name)
            ok = layers[int(m.group(1))].set_dependency(m.group(2), t) if m else glob.set_dependency(name, t)
            if not ok:
                unmapped.append(name)
        for i, c in enumerate(layers):
            if not c.is_populated:
                raise RuntimeError(f"layer {i}: checkpoint did not provide {c.unfinished()}")
        if not glob.is_populated:
            raise RuntimeError(f"checkpoint did not provide {glob.unfinished()}")
        weights: Dict[str, Any] = dict(glob.to_layer_weights())
        if "unembed" not in weights:
            weights["unembed"] = weights["embed"]              # tied embeddings
        weights["layers"] = [c.to_layer_weights() for c in layers]
        self.container_map.set_unmapped_params(unmapped)
        self._weights = weights
        return weights
