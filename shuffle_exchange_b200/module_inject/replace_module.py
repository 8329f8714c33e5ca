"""Model surgery entry points for inference (reference: ``deepspeed/module_inject/replace_module.py:182-690``).

``replace_transformer_layer`` in the reference either injects per-architecture fused kernels (containers) or runs
AutoTP; in this framework the fused-kernel path is the ragged v2 engine (``inference/v2``) and the v1 entry point keeps
the AutoTP behaviour: shard every block over the tensor-parallel group, slice ``lm_head``, optionally load sharded
checkpoints."""
from __future__ import annotations

import torch
import torch.nn as nn

from .. import comm as dist
from ..utils.logging import logger
from .auto_tp import AutoTP, Loading, ReplaceWithTensorSlicing
from .tp_shard import set_n_embd, set_num_attention_heads, set_num_kv_heads


def _module_match(module):
    from .replace_policy import generic_policies
    for policy in generic_policies:
        p = policy()
        if p.match(module):
            return p
    return None


def generic_injection(module, dtype=None, enable_cuda_graph=True):
    """Diffusers-style generic injection (attention blocks of UNet/VAE): kept as a no-op pass-through that reports what
    it would replace; the reference swaps in ``DeepSpeedDiffusersAttention`` (triton/CUDA kernels absent here)."""
    n = sum(1 for m in module.modules() if type(m).__name__ in ("CrossAttention", "Attention", "BasicTransformerBlock"))
    if n:
        logger.info(f"generic_injection: {n} diffusers attention blocks found; running them unmodified")
    return module


def replace_transformer_layer(orig_layer_impl, model, checkpoint_dict, config, model_config):
    """Apply AutoTP to ``model`` according to ``config.tensor_parallel`` (``tp_size``, ``tp_group``)."""
    tp = getattr(config, "tensor_parallel", None)
    mp_size = getattr(tp, "tp_size", 1) if tp is not None else 1
    mp_group = getattr(tp, "tp_group", None) if tp is not None else None
    if model_config is not None:
        kv = next((getattr(model_config, a) for a in ("num_key_value_heads", "num_kv_heads", "num_attention_heads")
                   if getattr(model_config, a, None)), None)
        set_num_kv_heads(kv)
        set_num_attention_heads(getattr(model_config, "num_attention_heads", None))
        set_n_embd(getattr(model_config, "hidden_size", None))
    if mp_size <= 1:
        return model
    policies = config.injection_policy_tuple if getattr(config, "injection_policy_tuple", None) else None
    parsed = AutoTP.tp_parser(model) if policies is None else [(orig_layer_impl, list(policies))]
    for block_cls, gem in parsed:
        worker = AutoTP(model, gem, "", checkpoint_dict, None, block_cls,
                        keep_module_on_host=getattr(config, "keep_module_on_host", False))
        worker.set_tensor_parallel_config(mp_size, mp_group)
        worker.update_policy_list()
        for m in model.modules():
            if type(m) is block_cls:
                worker.update_mp_params(m)
                worker._replace_module(m)
        worker._replace_last_linear_module(model)
    return model


def revert_transformer_layer(orig_layer_impl, model, config, preln=False):
    raise NotImplementedError("reverting kernel injection is only meaningful for container-based injection, which this "
                              "framework replaces with the ragged v2 engine")


def replace_module(model, orig_class, replace_fn, _replace_policy, checkpoint=None):
    """Generic walker: replace every instance of ``orig_class`` by ``replace_fn(child, policy, layer_id)``."""
    layer_id = 0
    for name, child in list(model.named_children()):
        if isinstance(child, orig_class):
            setattr(model, name, replace_fn(child, _replace_policy, layer_id))
            layer_id += 1
        else:
            replace_module(child, orig_class, replace_fn, _replace_policy, checkpoint)
    return model
