"""Swap HuggingFace BERT layers for the fused *training* transformer layer (reference: ``module_inject/inject.py``).

``module_inject(BertLayer, model, config, micro_batch_size, max_seq_length, seed, preln, fp16)`` walks ``model`` and
replaces every ``layer_obj`` instance by a :class:`DeepSpeedTransformerLayer` initialised from the client's weights
(q, k, v fused into one ``[3H, H]`` projection), so fine-tuning continues from the same function."""
from __future__ import annotations

import copy

import torch

from ..ops.transformer import DeepSpeedTransformerConfig, DeepSpeedTransformerLayer


def _bert_layer_to_fused(child, config, micro_batch_size, seed, preln, fp16, bf16=False):
    cfg = copy.copy(config)
    hidden = child.attention.self.query.weight.shape[1]
    ds_cfg = DeepSpeedTransformerConfig(
        batch_size=micro_batch_size, hidden_size=hidden, intermediate_size=child.intermediate.dense.out_features,
        heads=child.attention.self.num_attention_heads,
        attn_dropout_ratio=getattr(cfg, "attention_probs_dropout_prob", 0.0),
        hidden_dropout_ratio=getattr(cfg, "hidden_dropout_prob", 0.0),
        num_hidden_layers=getattr(cfg, "num_hidden_layers", -1), initializer_range=getattr(cfg, "initializer_range", 0.02),
        layer_norm_eps=getattr(cfg, "layer_norm_eps", 1e-12), seed=seed, fp16=fp16, bf16=bf16, pre_layer_norm=preln,
        local_rank=-1, training=True)
    new = DeepSpeedTransformerLayer(ds_cfg)
    s = child.attention.self
    with torch.no_grad():
        new.attn_qkvw.copy_(torch.cat([s.query.weight, s.key.weight, s.value.weight], dim=0))
        new.attn_qkvb.copy_(torch.cat([s.query.bias, s.key.bias, s.value.bias], dim=0))
        new.attn_ow.copy_(child.attention.output.dense.weight)
        new.attn_ob.copy_(child.attention.output.dense.bias)
        # post-LN BERT: attention.output.LayerNorm follows attention, output.LayerNorm ends the layer; a pre-LN layer
        # (``PreAttentionLayerNorm`` / ``PostAttentionLayerNorm`` attributes) maps the same way the reference does
        attn_ln = getattr(child, "PostAttentionLayerNorm", None) if preln else child.attention.output.LayerNorm
        final_ln = getattr(child, "PreAttentionLayerNorm", None) if preln else child.output.LayerNorm
        attn_ln = attn_ln if attn_ln is not None else child.attention.output.LayerNorm
        final_ln = final_ln if final_ln is not None else child.output.LayerNorm
        new.attn_nw.copy_(attn_ln.weight)
        new.attn_nb.copy_(attn_ln.bias)
        new.inter_w.copy_(child.intermediate.dense.weight)
        new.inter_b.copy_(child.intermediate.dense.bias)
        new.output_w.copy_(child.output.dense.weight)
        new.output_b.copy_(child.output.dense.bias)
        new.norm_w.copy_(final_ln.weight)
        new.norm_b.copy_(final_ln.bias)
    ref = child.intermediate.dense.weight
    return new.to(device=ref.device, dtype=ref.dtype)


def module_inject(layer_obj, model, config, micro_batch_size, max_seq_length, seed, preln, fp16=True, bf16=False):
    for name, child in list(model.named_children()):
        if isinstance(child, layer_obj):
            setattr(model, name, _bert_layer_to_fused(child, config, micro_batch_size, seed, preln, fp16, bf16))
        else:
            module_inject(layer_obj, child, config, micro_batch_size, max_seq_length, seed, preln, fp16, bf16)
    return model
