"""AutoTP: shard an arbitrary HuggingFace-style transformer over a tensor-parallel group without per-architecture
containers.

Reference: ``deepspeed/module_inject/auto_tp.py:193-560`` — ``tp_parser`` walks the repeated transformer block and
decides, per linear layer, column-parallel (``LinearLayer``) vs row-parallel with all-reduce (``LinearAllreduce``);
``_replace_module`` swaps them in place, slices embeddings' consumers (``lm_head`` -> ``LmHeadLinearAllreduce``) and
rescales head-count attributes of the attention modules (``update_mp_params``).

The classification here is structural first (the LAST linear of every attention / MLP sub-block reduces), with the
reference's name lists as tie-breakers, so custom blocks (e.g. this framework's fused ``wqkv`` / ``w_gate_up`` layout
exposed through ``nn.Linear``) parse without a policy.
"""
from __future__ import annotations

import re
from typing import Dict, List, Optional, Tuple

import torch
import torch.nn as nn

from .. import comm as dist
from ..utils.logging import logger
from .layers import (Conv_LinearALlreduce, GateUpPack_LinearLayer, LinearAllreduce, LinearLayer,
                     LmHeadLinearAllreduce, conv_LinearLayer, fused_LinearLayer)
from .tp_shard import get_shard_size, get_shard_size_list, set_n_embd, set_num_attention_heads, set_num_kv_heads

ROW_PARALLEL_NAMES = ("o_proj", "out_proj", "down_proj", "dense_4h_to_h", "attention.dense", "self_attention.dense",
                      "attn.c_proj", "mlp.c_proj", "fc2", "w2", "wo", "w_down", "attn.out_proj", "proj_out",
                      "attention.output.dense", "output.dense", "final_linear", "ff_out", "attn_out")
FUSED_QKV_NAMES = ("query_key_value", "qkv_proj", "c_attn", "wqkv", "Wqkv", "qkv", "att_proj")
GATE_UP_NAMES = ("gate_up_proj", "w_gate_up", "gate_up")
HEAD_ATTRS = ("num_heads", "num_attention_heads", "n_head", "n_heads", "num_attn_heads", "num_kv_heads",
              "num_key_value_heads", "num_kv", "all_head_size", "embed_dim", "hidden_size", "split_size",
              "num_key_value_groups", "head_size")


class ReplaceWithTensorSlicing:
    """Copy a slice of a full tensor into a (smaller) destination according to the TP rank (reference :31-190)."""

    def __init__(self, mp_group=None, mp_size=1, out_dim=1, in_dim=0):
        self.gpu_index = dist.get_rank(group=mp_group) if (mp_group is not None and dist.is_initialized()) else 0
        self.out_dim, self.in_dim, self.mp_size = out_dim, in_dim, mp_size

    def merge_assert(self, dim1, dim2):
        assert dim1 > dim2, "Merging tensors is not allowed here! Please use deepspeed load_checkpoint for merging your checkpoints before replacing the transformer layer with inference-kernels"

    def strided_copy(self, dst, src, num_splits: int, int8: bool = False, allocate_tensor: bool = False):
        """``src`` is a fusion of ``num_splits`` equal blocks along its output dim; take this rank's slice of each."""
        if src is None:
            return src
        src_shape, dst_shape = src.shape, dst.shape
        outer = 0 if len(src_shape) == 1 else (self.out_dim if not int8 else self.in_dim)
        if src_shape == dst_shape or src_shape[outer] == dst_shape[outer if len(dst_shape) > outer else 0]:
            return src
        blocks = torch.chunk(src, num_splits, dim=outer)
        mine = [torch.chunk(b, self.mp_size, dim=outer)[self.gpu_index] for b in blocks]
        out = torch.cat(mine, dim=outer)
        if allocate_tensor:
            return out.clone()
        dst.data.copy_(out.reshape(dst.shape))
        return dst

    def copy(self, dst, src, int8: bool = False, allocate_tensor: bool = False):
        if src is None:
            return src
        if src.shape == dst.shape:
            dst.data.copy_(src)
            return dst
        for dim in range(src.dim()):
            if src.shape[dim] != dst.shape[dim]:
                self.merge_assert(src.shape[dim], dst.shape[dim])
                piece = torch.chunk(src, self.mp_size, dim=dim)[self.gpu_index]
                if allocate_tensor:
                    return piece.clone()
                dst.data.copy_(piece)
                return dst
        return dst


def get_transformer_name(replaced_module) -> str:
    from torch.nn import ModuleList
    for n, c in replaced_module.named_children():
        for cn, cc in c.named_children():
            if isinstance(cc, ModuleList):
                return n + "." + cn
        if isinstance(c, ModuleList):
            return n
    return ""


class Loading:
    @staticmethod
    def is_load_module(module):
        return isinstance(module, (nn.Linear, nn.Embedding, nn.LayerNorm)) or type(module).__name__ in (
            "LlamaRMSNorm", "MistralRMSNorm", "Qwen2RMSNorm", "T5LayerNorm", "RMSNorm", "LPLayerNorm")


class AutoTP:
    def __init__(self, module, all_reduce_linears, prefix, state_dict, linear_layer_setting, orig_layer_impl,
                 keep_module_on_host=False):
        self.module = module
        self.all_reduce_linears = all_reduce_linears
        self.prefix = prefix
        self.state_dict = state_dict
        self.mp_size = None
        self.mp_group = None
        self.linear_layer_setting = linear_layer_setting
        self.orig_layer_impl = orig_layer_impl
        self.linear_policies = None
        self.conv_linear_layer = False
        self.keep_module_on_host = keep_module_on_host
        from .layers import TensorParallel_Layer
        TensorParallel_Layer.set_keep_module_on_host(keep_module_on_host)

    # ------------------------------------------------------------------------------------- parsing
    @staticmethod
    def in_module_list(module, module_list):
        return any(type(item).__name__ == type(module).__name__ for item in module_list)

    @staticmethod
    def get_module_list(model) -> List[nn.Module]:
        """One representative of every distinct block class found inside ``nn.ModuleList`` containers."""
        out = []
        for child in model.children():
            if isinstance(child, nn.ModuleList):
                for m in child.children():
                    if not out or not AutoTP.in_module_list(m, out):
                        out.append(m)
            else:
                out += [m for m in AutoTP.get_module_list(child) if not AutoTP.in_module_list(m, out)]
        return out

    @staticmethod
    def supported(model) -> bool:
        unsupported = ("deberta", "flaubert", "fsmt", "gpt2", "led", "longformer", "xlm", "xlnet")
        name = str(type(model)).lower()
        return not any(re.search(rf"\b{u}\b|{u}", name) for u in unsupported if u != "gpt2") or "gpt2" not in name

    @staticmethod
    def _linear_children(block: nn.Module) -> List[Tuple[str, nn.Module]]:
        return [(n, m) for n, m in block.named_modules() if isinstance(m, nn.Linear) or type(m).__name__ == "Conv1D"]

    @staticmethod
    def tp_parser(model) -> List[Tuple[type, List[str]]]:
        """[(block class, [names of the linears whose output must be all-reduced])] — structural rule: inside every
        direct sub-module of the block that owns >= 2 linears (attention, MLP), the last linear reduces; known names
        (``ROW_PARALLEL_NAMES``) are added regardless."""
        policy = []
        for block in AutoTP.get_module_list(model):
            gem: List[str] = []
            for sub_name, sub in block.named_children():
                lin = AutoTP._linear_children(sub)
                if len(lin) >= 2:
                    gem.append(f"{sub_name}.{lin[-1][0]}".split(".")[-1] if False else lin[-1][0].split(".")[-1])
            for name, _ in AutoTP._linear_children(block):
                if any(name.endswith(k) or k in name for k in ROW_PARALLEL_NAMES):
                    gem.append(name.split(".")[-1])
            gem = sorted(set(gem))
            assert gem, f"AutoTP could not find the row-parallel linears of {type(block).__name__}; pass an injection policy"
            policy.append((type(block), gem))
        return policy

    # --------------------------------------------------------------------------------- replacement
    def set_tensor_parallel_config(self, mp_size, mp_group):
        self.mp_size, self.mp_group = mp_size, mp_group

    def update_policy_list(self):
        self.linear_policies = {nn.Linear: self._replace, nn.Embedding: self._slice_embedding}

    def _replace(self, child, name, conv_linear_layer=False):
        if getattr(child, "replaced", False):
            return child
        is_conv = type(child).__name__ == "Conv1D" or conv_linear_layer
        short = name.split(".")[-1]
        kw = dict(name=name)
        if short in self.all_reduce_linears or any(name.endswith(k) for k in self.all_reduce_linears):
            if is_conv:
                return Conv_LinearALlreduce(child, self.mp_group, **kw)
            if short in ("lm_head", "embed_out"):
                return LmHeadLinearAllreduce(child, self.mp_group, **kw)
            return LinearAllreduce(child, self.mp_group, **kw)
        if any(k in name for k in GATE_UP_NAMES):
            return GateUpPack_LinearLayer(child, self.mp_group, **kw)
        if any(k in name for k in FUSED_QKV_NAMES):
            return fused_LinearLayer(child, self.mp_group, fused_parts=3, **kw) if not is_conv else conv_LinearLayer(child, self.mp_group, **kw)
        if is_conv:
            return conv_LinearLayer(child, self.mp_group, **kw)
        return LinearLayer(child, self.mp_group, **kw)

    def _slice_embedding(self, child, name, conv_linear_layer=False):
        return child            # embeddings stay replicated (the reference slices only when mp_replace is forced)

    def update_mp_params(self, child):
        if getattr(child, "replaced", False):
            return
        for attr in HEAD_ATTRS:
            if hasattr(child, attr) and isinstance(getattr(child, attr), int):
                val = getattr(child, attr)
                if attr in ("num_key_value_groups", "head_size"):
                    continue
                setattr(child, attr, get_shard_size(val, self.mp_size))
        setattr(child, "replaced", True)

    def _replace_module(self, r_module, prev_name="", prev_class_name=""):
        for name, child in list(r_module.named_children()):
            full = name if prev_class_name == "" else prev_class_name + "." + name
            if isinstance(child, nn.Linear) or type(child).__name__ == "Conv1D":
                new = self._replace(child, full, self.conv_linear_layer)
                setattr(r_module, name, new)
            elif isinstance(child, nn.Embedding):
                setattr(r_module, name, self._slice_embedding(child, full))
            else:
                self.update_mp_params(child)
                self._replace_module(child, name, full)
        return r_module

    def get_model_num_kv_heads(self, config):
        for attr in ("num_kv_heads", "num_key_value_heads", "num_attention_heads", "n_heads", "attention_heads"):
            v = getattr(config, attr, None)
            if v is not None:
                return v
        return None

    def _replace_last_linear_module(self, r_module):
        if hasattr(r_module, "lm_head") and isinstance(r_module.lm_head, nn.Linear):
            setattr(r_module, "lm_head", LmHeadLinearAllreduce(r_module.lm_head, self.mp_group, name="lm_head"))
        elif hasattr(r_module, "embed_out") and isinstance(r_module.embed_out, nn.Linear):
            setattr(r_module, "embed_out", LmHeadLinearAllreduce(r_module.embed_out, self.mp_group, name="embed_out"))
        return r_module


def tp_model_init(model, tp_size: int, dtype, config=None, tp_group=None, **kwargs):
    """Shard ``model`` in place over ``tp_group`` (training entry point; reference ``deepspeed.tp_model_init``)."""
    from .layers import set_autotp_mode
    set_autotp_mode(training=True)
    if tp_group is None and dist.is_initialized() and tp_size > 1:
        world, rank = dist.get_world_size(), dist.get_rank()
        for start in range(0, world, tp_size):
            g = dist.new_group(list(range(start, start + tp_size)))
            if start <= rank < start + tp_size:
                tp_group = g
    cfg = getattr(model, "config", None)
    if cfg is not None:
        kv = next((getattr(cfg, a) for a in ("num_key_value_heads", "num_kv_heads", "num_attention_heads") if getattr(cfg, a, None)), None)
        set_num_kv_heads(kv)
        set_num_attention_heads(getattr(cfg, "num_attention_heads", None))
        set_n_embd(getattr(cfg, "hidden_size", None))
    policies = AutoTP.tp_parser(model)
    for block_cls, gem in policies:
        tp = AutoTP(model, gem, "", None, None, block_cls)
        tp.set_tensor_parallel_config(tp_size, tp_group)
        tp.update_policy_list()
        for m in model.modules():
            if type(m) is block_cls:
                tp.update_mp_params(m)
                tp._replace_module(m)
    model.to(dtype) if dtype is not None else None
    model.ds_autotp_parsed = True
    return model
