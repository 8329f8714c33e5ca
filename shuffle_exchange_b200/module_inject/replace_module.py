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


class GroupQuantizer:
    """Symmetric group-wise integer quantiser of injected GEMM weights (reference ``replace_module.py:44-77``): the
    rows of a weight are cut into ``num_groups`` (or ``rows / group_size``) consecutive groups, each with one step size
    ``2 * absmax / 2**bits``; the result is an int8 ``Parameter`` carrying ``.scale`` (fp32, one entry per group) that
    ``ops.transformer.inference.op_binding.dequantize_int8`` expands at use.  With ``q_int8=False`` (or a tensor that is
    not to be quantised) the weight is returned as a frozen ``Parameter`` with an empty ``.scale``, as the callers of
    the reference class expect."""

    def __init__(self, q_int8: bool = True, group_size: int = 1, num_bits: int = 8, num_groups: int = 0):
        self.q_int8, self.group_size, self.num_bits, self.num_groups = q_int8, group_size, num_bits, num_groups

    def groups_for(self, rows: int) -> int:
        g = self.num_groups if self.num_groups > 0 else max(1, rows // max(1, self.group_size))
        while rows % g:                     # the groups must tile the rows
            g -= 1
        return g

    def quantize(self, inputs: torch.Tensor, qkv: bool = True, count: int = 1, parallel_dim: int = 0):
        if not self.q_int8 or not qkv or inputs.dim() != 2:
            out = nn.Parameter(inputs, requires_grad=False)
            out.scale = torch.empty(1)
            return out
        levels = 1 << self.num_bits
        g = self.groups_for(inputs.shape[0])
        flat = inputs.detach().float().reshape(g, -1)
        scale = flat.abs().amax(dim=1, keepdim=True).clamp_min(1e-12) * 2.0 / levels
        q = (flat / scale).round_().clamp_(-(levels // 2), levels // 2 - 1)
        out = nn.Parameter(q.reshape(inputs.shape).to(torch.int8).contiguous(), requires_grad=False)
        out.scale = scale.reshape(-1).contiguous()
        return out


def _module_match(module):
    from .replace_policy import generic_policies
    for policy in generic_policies:
        p = policy()
        if p.match(module):
            return p
    return None


def generic_injection(module, dtype=None, enable_cuda_graph=True):
    """Diffusers pipelines (reference ``replace_module.py:86-180``): wrap ``unet`` / ``vae`` / ``text_encoder``
    sub-models in their graphed wrappers; a bare UNet / VAE is wrapped directly."""
    for policy in (p() for p in __import__(__package__ + ".replace_policy", fromlist=["generic_policies"]).generic_policies):
        if policy.match(module):
            return policy.apply(module, enable_cuda_graph=enable_cuda_graph)
    wrapped = 0
    for name in ("unet", "vae", "text_encoder"):
        sub = getattr(module, name, None)
        if sub is None:
            continue
        p = _module_match(sub)
        if p is not None:
            setattr(module, name, p.apply(sub, enable_cuda_graph=enable_cuda_graph))
            wrapped += 1
        elif name == "text_encoder" and type(sub).__name__ == "CLIPTextModel":
            from ..model_implementations.transformers.clip_encoder import DSClipEncoder
            setattr(module, name, DSClipEncoder(sub, enable_cuda_graph=enable_cuda_graph))
            wrapped += 1
    if wrapped:
        logger.info(f"generic_injection: wrapped {wrapped} diffusers sub-model(s)")
    return module


def _policy_for_layer(module, injection_policy=None):
    from .containers import POLICY_TO_CONTAINER
    from .replace_policy import replace_policies
    for pol in (injection_policy,) if injection_policy is not None else replace_policies:
        if pol in POLICY_TO_CONTAINER and pol.match(module):
            return pol
    return None


def _inject_kernels(model, config, model_config, injection_policy=None):
    """``replace_with_kernel_inject``: every layer a policy recognises becomes a fused
    :class:`DeepSpeedTransformerInference` carrying the client's weights (tensor-parallel slices when ``tp_size > 1``);
    returns the number of layers replaced (reference ``replace_module.py:261-411``)."""
    from .containers import POLICY_TO_CONTAINER, HFLayerAdapter
    from ..model_implementations.transformers.ds_transformer import DeepSpeedTransformerInference
    from ..ops.transformer.inference.ds_attention import DeepSpeedSelfAttention
    DeepSpeedTransformerInference.layer_id = 0
    DeepSpeedSelfAttention.num_layers = 0
    count = 0

    def walk(parent):
        nonlocal count
        for name, child in list(parent.named_children()):
            pol = _policy_for_layer(child, injection_policy)
            if pol is None:
                walk(child)
                continue
            policy = pol(child, inference=True)
            container = POLICY_TO_CONTAINER[pol](policy, config, model_config, count, child)
            container.create_ds_model_config()
            container.initialize_tensors()
            container.apply_tensor_parallelism()
            fused = container.create_module()
            dev = next(child.parameters()).device
            adapter = HFLayerAdapter(fused.to(dev), returns_tuple=pol.returns_tuple)
            adapter._orig_layer = [child]          # kept out of the module tree so revert can restore it
            setattr(parent, name, adapter)
            count += 1

    walk(model)
    return count


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
    if getattr(config, "replace_with_kernel_inject", False):
        n = _inject_kernels(model, config, model_config, orig_layer_impl if isinstance(orig_layer_impl, type) and
                            hasattr(orig_layer_impl, "get_hidden_heads") else None)
        if n == 0:
            raise ValueError("replace_with_kernel_inject: no layer of this model matches an injection policy "
                             "(supported: BERT/RoBERTa, DistilBERT, GPT-2, GPT-J, GPT-Neo, GPT-NeoX, BLOOM, OPT, LLaMA, "
                             "InternLM, CLIP, Megatron-GPT); use AutoTP (kernel_inject=False) instead")
        logger.info(f"kernel injection: {n} transformer layers replaced")
        return model
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
    """Put the client layers back where :func:`replace_transformer_layer` injected fused ones (weights untouched:
    injection copies, it does not move)."""
    from .containers import HFLayerAdapter
    for parent in list(model.modules()):
        for name, child in list(parent.named_children()):
            if isinstance(child, HFLayerAdapter) and getattr(child, "_orig_layer", None):
                setattr(parent, name, child._orig_layer[0])
    return model


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
