"""Load Megatron-style tensor-parallel checkpoints at a different TP degree (merge or split on the fly).

Reference: ``deepspeed/runtime/state_dict_factory.py:21-427`` — ``SDLoaderFactory.get_sd_loader[_json]``,
``SDLoaderBase.load`` (same-degree load, merge when there are more files than ranks, split when fewer) and
``MegatronSDLoader`` with the three historical fused-QKV layouts (checkpoint version 0 / 1.0 / 2.0), optional INT8
weight quantisation while loading.  The per-key behaviour lives in one rule table here (``_RULES``) that both the merge
and the split path consult."""
from __future__ import annotations

import collections
import copy
import json
import os
from abc import ABC, abstractmethod
from typing import List, Optional, Tuple

import torch

from ..utils.logging import logger
from .checkpoint_engine.torch_checkpoint_engine import TorchCheckpointEngine
from .weight_quantizer import WeightQuantization

AUTO_MODULE_KEY = "auto"

# (substring of the key, tensor axis that is partitioned across TP ranks, kind)
#   kind "row": row-parallel weight (axis 1)          kind "col": column-parallel weight / bias / embedding (axis 0)
#   kind "qkv": fused QKV, axis 0 with a version-dependent interleave
_RULES: Tuple[Tuple[str, int, str], ...] = (
    ("attention.dense.weight", 1, "row"),
    ("mlp.dense_4h_to_h.weight", 1, "row"),
    ("attention.query_key_value", 0, "qkv"),
    ("mlp.dense_h_to_4h.weight", 0, "col"),
    ("mlp.dense_h_to_4h.bias", 0, "col"),
    ("word_embeddings.weight", 0, "col"),
    ("final_linear.weight", 0, "col"),
)
_QUANTIZED = ("attention.dense.weight", "mlp.dense_4h_to_h.weight", "attention.query_key_value.weight",
              "mlp.dense_h_to_4h.weight")


def _rule(key: str, for_merge: bool) -> Optional[Tuple[int, str]]:
    for pat, axis, kind in _RULES:
        if pat in key:
            if for_merge and pat == "final_linear.weight":     # the reference only splits this one
                return None
            return axis, kind
    return None


class SDLoaderFactory:

    @staticmethod
    def get_sd_loader_json(json_file, checkpoint_engine):
        if isinstance(json_file, str):
            with open(json_file) as f:
                data = json.load(f)
        else:
            assert isinstance(json_file, dict)
            data = json_file
        sd_type = data["type"]
        if sd_type.lower() in ("bloom", "ds_model"):
            return data
        return SDLoaderFactory.get_sd_loader(data["checkpoints"], checkpoint_engine, sd_type, data["version"])

    @staticmethod
    def get_sd_loader(ckpt_list, checkpoint_engine, sd_type="Megatron", version=None):
        if sd_type == "Megatron":
            return MegatronSDLoader(ckpt_list, version, checkpoint_engine)
        raise AssertionError(f"{sd_type} checkpoint type is not supported")


class SDLoaderBase(ABC):

    def __init__(self, ckpt_list, version, checkpoint_engine):
        self.module_key = None
        self.ckpt_list = ckpt_list
        self.version = version
        self.checkpoint_engine = TorchCheckpointEngine() if checkpoint_engine is None else checkpoint_engine
        self.check_ckpt_list()

    def _load_file(self, path):
        return self.checkpoint_engine.load(path, map_location=lambda storage, loc: storage)

    def load(self, mp_world_size, mp_rank, module_key=AUTO_MODULE_KEY, is_pipe_parallel=False, quantize=False,
             quantize_bits=8, quantize_groups=64, mlp_extra_grouping=True):
        """-> ``(load_path, state_dict, (scales_or_None, merge_count))``.  Three cases: as many files as ranks (load
        own file), more files than ranks (merge ``n/ranks`` files), fewer (split one file ``ranks/n`` ways)."""
        self.module_key = module_key
        num_ckpt = len(self.ckpt_list)
        idx = mp_rank * num_ckpt // mp_world_size
        if is_pipe_parallel and module_key is not None and mp_world_size != num_ckpt:
            # mp_rank_* files of a pipeline module all hold the same content when the degree changed: take the first
            mp_world_size, idx = num_ckpt, 0
        load_path = self.ckpt_list[idx]
        merge_count = 1
        if num_ckpt == mp_world_size:
            assert os.path.exists(load_path)
            sd = self._load_file(load_path)
            all_scales = None
            if quantize:
                quantizer = WeightQuantization(mlp_extra_grouping=mlp_extra_grouping, mp_size=mp_world_size)
                module, all_scales = quantizer.sd_quantize_megatron(self.get_module(sd), quantize_bits,
                                                                    quantize_groups)
                self.set_module(sd, module)
        elif num_ckpt > mp_world_size:
            sd, all_scales, merge_count = self.merge_state_dict(mp_world_size, mp_rank, quantize, quantize_bits,
                                                                quantize_groups, mlp_extra_grouping)
        else:
            sd, all_scales = self.split_state_dict(mp_world_size, mp_rank, quantize, quantize_bits, quantize_groups,
                                                   mlp_extra_grouping)
        return load_path, sd, (all_scales, merge_count)

    def get_merge_state_dicts(self, mp_world_size, mp_rank):
        num_ckpt = len(self.ckpt_list)
        assert num_ckpt % mp_world_size == 0, "Invalid checkpoints and world size for sd merge"
        n = num_ckpt // mp_world_size
        mine = self.ckpt_list[n * mp_rank:n * (mp_rank + 1)]
        logger.info(f"mp_rank: {mp_rank}, ckpt_list: {mine}")
        return [self._load_file(c) for c in mine]

    def get_split_state_dict(self, mp_world_size, mp_rank):
        num_ckpt = len(self.ckpt_list)
        assert mp_world_size % num_ckpt == 0, "Invalid checkpoints and world size for sd split"
        num_to_split = mp_world_size // num_ckpt
        index, offset = divmod(mp_rank, num_to_split)
        logger.info(f"mp_rank: {mp_rank}, ckpt_list: {self.ckpt_list[index]}, offset: {offset}")
        return self._load_file(self.ckpt_list[index]), num_to_split, offset

    def _choose_module_key(self, sd):
        assert not ("module" in sd and "model" in sd), \
            "checkpoint has both 'model' and 'module' keys, not sure how to proceed"
        assert "module" in sd or "model" in sd, \
            "checkpoint contains neither 'model' or 'module' keys, not sure how to proceed"
        return "module" if "module" in sd else "model"

    def get_module(self, sd):
        if self.module_key is None:
            return sd
        if self.module_key == AUTO_MODULE_KEY:
            return sd[self._choose_module_key(sd)]
        return sd[self.module_key]

    def set_module(self, sd, module):
        if self.module_key is None:
            return module
        if self.module_key == AUTO_MODULE_KEY:
            sd[self._choose_module_key(sd)] = module
        else:
            sd[self.module_key] = module
        return sd

    def check_ckpt_list(self):
        assert len(self.ckpt_list) > 0
        sd = self._load_file(self.ckpt_list[0])
        if "mp_world_size" in sd.keys():
            assert len(self.ckpt_list) == sd["mp_world_size"], \
                f"checkpoint count {len(self.ckpt_list)} is different from saved mp_world_size {sd['mp_world_size']}"

    @abstractmethod
    def merge_state_dict(self, mp_world_size, mp_rank, quantize, quantize_bits, groups, mlp_extra_grouping):
        ...

    @abstractmethod
    def split_state_dict(self, mp_world_size, mp_rank, quantize, quantize_bits, groups, mlp_extra_grouping):
        ...

    @abstractmethod
    def sanity_check(self, ckpt_file_name):
        ...


class MegatronSDLoader(SDLoaderBase):
    """Fused-QKV layouts by checkpoint version (``h`` hidden, ``n`` heads, ``p`` TP degree, ``hn = h/n``):
    v0 ``[(3 * n/p * hn), h]`` — Q, K, V blocks per rank (must be regrouped when the degree changes);
    v1.0 ``[(n/p * hn * 3), h]`` and v2.0 ``[(n/p * 3 * hn), h]`` — head-major, plain concatenation/splitting."""

    def merge_query_key_value(self, param_list, ckpt_ver):
        if ckpt_ver == 0:
            assert param_list[0].shape[0] % 3 == 0
            third = param_list[0].shape[0] // 3
            pieces = [torch.split(p, third, dim=0) for p in param_list]
            return torch.cat([torch.cat([pc[i] for pc in pieces], dim=0) for i in range(3)], dim=0)
        if ckpt_ver in (1.0, 2.0):
            return torch.cat(param_list, dim=0)
        raise AssertionError(f"checkpoint version: {ckpt_ver} is not supported")

    def split_query_key_value(self, param, num_to_split, offset, ckpt_ver):
        if ckpt_ver == 0:
            assert param.shape[0] % 3 == 0
            q, k, v = torch.split(param, param.shape[0] // 3, dim=0)
            assert q.shape[0] % num_to_split == 0
            size = q.shape[0] // num_to_split
            return torch.cat([torch.split(t, size, dim=0)[offset] for t in (q, k, v)], dim=0)
        if ckpt_ver in (1.0, 2.0):
            assert param.shape[0] % num_to_split == 0
            return torch.split(param, param.shape[0] // num_to_split, dim=0)[offset]
        raise AssertionError(f"checkpoint version: {ckpt_ver} is not supported")

    def merge_state_dict(self, mp_world_size, mp_rank, quantize=False, quantize_bits=8, groups=64,
                         mlp_extra_grouping=True):
        self.sanity_check(self.ckpt_list[0])
        sd_list = self.get_merge_state_dicts(mp_world_size, mp_rank)
        ds_sd = copy.deepcopy(sd_list[0])
        modules = [self.get_module(sd) for sd in sd_list]
        ver = self.get_checkpoint_version(ds_sd)
        logger.info(f"checkpoint version: {ver}")
        quantizer = WeightQuantization(mlp_extra_grouping=mlp_extra_grouping, mp_size=mp_world_size) \
            if quantize else None
        merged = collections.OrderedDict()
        for key in modules[0].keys():
            values = [m[key] for m in modules]
            rule = _rule(key, for_merge=True)
            if rule is None:
                merged[key] = values[0]
                continue
            axis, kind = rule
            do_q = quantize and any(q in key for q in _QUANTIZED)
            if do_q:
                values = quantizer.Quantize(values, quantize_bits, groups, key=key, merge_dim=axis)
            if kind == "qkv" and not quantize:
                merged[key] = self.merge_query_key_value(values, ver)
            else:
                merged[key] = torch.cat(values, dim=axis)
        ds_sd = self.set_module(ds_sd, merged)
        return ds_sd, (quantizer.merge_scales() if quantize else None), len(modules)

    def split_state_dict(self, mp_world_size, mp_rank, quantize=False, quantize_bits=8, groups=64,
                         mlp_extra_grouping=True):
        sd, num_to_split, offset = self.get_split_state_dict(mp_world_size, mp_rank)
        ds_sd = copy.deepcopy(sd)
        module = self.get_module(sd)
        ver = self.get_checkpoint_version(ds_sd)
        logger.info(f"checkpoint version: {ver}")
        quantizer = WeightQuantization(mlp_extra_grouping=mlp_extra_grouping, mp_size=mp_world_size) \
            if quantize else None
        out = collections.OrderedDict()
        for key, value in module.items():
            rule = _rule(key, for_merge=False)
            if rule is None:
                out[key] = value
                continue
            axis, kind = rule
            if quantize and any(q in key for q in _QUANTIZED):
                value = quantizer.Quantize([value], quantize_bits, groups, key)[0]
            if kind == "qkv":
                out[key] = self.split_query_key_value(value, num_to_split, offset, ver)
            else:
                assert value.shape[axis] % num_to_split == 0
                out[key] = torch.split(value, value.shape[axis] // num_to_split, dim=axis)[offset]
        ds_sd = self.set_module(ds_sd, out)
        return ds_sd, (quantizer.merge_scales_split(num_to_split) if quantize else None)

    def sanity_check(self, ckpt_file_name):
        needed = ["attention.dense.weight", "mlp.dense_4h_to_h.weight", "attention.query_key_value",
                  "mlp.dense_h_to_4h.weight", "mlp.dense_h_to_4h.bias"]
        keys = list(self.get_module(self._load_file(ckpt_file_name)).keys())
        for part in needed:
            assert any(part in k for k in keys), f"key: {part} is not found in the checkpoint {ckpt_file_name}"

    def get_checkpoint_version(self, state_dict):
        return self.version if self.version is not None else state_dict.get("checkpoint_version", 0)
