"""User entry points of the compression library (reference: ``deepspeed/compression/compress.py:20-235``)."""
from __future__ import annotations

import json
import os
import re

from ..utils.logging import logger
from .config import get_compression_config
from .constants import *  # noqa: F401,F403
from .helper import compression_preparation, fix_compression, is_module_compressible, recursive_getattr


def check_deepspeed_config(config):
    if isinstance(config, dict):
        return config
    if os.path.exists(config):
        return json.load(open(config, "r"), object_pairs_hook=dict)
    raise ValueError(f"Expected a string path to an existing deepspeed config, or a dictionary. Received: {config}")


def get_module_name(group_name, model, key_word, exist_module_name, mpu=None, verbose=True):
    """Names of compressible modules matching ``key_word`` (regex / substring; ``*`` = all) not yet claimed by a group."""
    picked = []
    for name, module in model.named_modules():
        if not is_module_compressible(module, mpu):
            continue
        if key_word == "*" or re.search(key_word, name) is not None:
            if name in exist_module_name and verbose:
                raise ValueError(f"{name} is already added to compression, please check your config file for {group_name}.")
            if name not in exist_module_name:
                exist_module_name.add(name)
                picked.append(name)
    return picked, exist_module_name


def get_compress_methods(model, compress_methods, mpu=None):
    """[[module names, related module names, {technique: params}], ...] for every enabled technique group."""
    out = []
    for method, content in compress_methods.items():
        if method == LAYER_REDUCTION:
            continue
        exist = set()
        shared = content[SHARED_PARAMETERS]
        for gname, g in content[DIFFERENT_GROUPS].items():
            names = []
            scopes = g[DIFFERENT_GROUPS_MODULE_SCOPE]
            for kw in (scopes if isinstance(scopes, (list, tuple)) else [scopes]):
                found, exist = get_module_name(gname, model, kw, exist, mpu=mpu)
                names.extend(found)
            related = []
            if g[DIFFERENT_GROUPS_RELATED_MODULE_SCOPE]:
                for kws in g[DIFFERENT_GROUPS_RELATED_MODULE_SCOPE]:
                    tmp = []
                    for kw in kws:
                        found, _ = get_module_name(gname, model, kw, set(), mpu=mpu)
                        tmp.append(found)
                    related.append(tmp)
            params = dict(shared)
            params.update(g[DIFFERENT_GROUPS_PARAMETERS])
            out.append([names, related, {method: params}])
    return out


def init_compression(model, deepspeed_config, teacher_model=None, mpu=None):
    """Replace compressible layers and arm the configured techniques (they start acting when the scheduler enables
    them at their ``schedule_offset``)."""
    cfg = get_compression_config(check_deepspeed_config(deepspeed_config))
    c_model = model.module if hasattr(model, "module") else model
    if cfg[LAYER_REDUCTION][LAYER_REDUCTION_ENABLED]:
        assert teacher_model is not None, "Teacher model is required for layer reduction"
        student_initialization(c_model, teacher_model, deepspeed_config)
    methods = get_compress_methods(c_model, cfg, mpu=mpu)
    compression_preparation(c_model, methods, mpu)
    return model


def redundancy_clean(model, deepspeed_config, mpu=None):
    """After training: bake masks and quantisation into the weights, shrinking pruned dimensions (and the dimensions of
    the ``related_modules`` that consume them)."""
    cfg = get_compression_config(check_deepspeed_config(deepspeed_config))
    c_model = model.module if hasattr(model, "module") else model
    methods = get_compress_methods(c_model, cfg, mpu=mpu)
    order = [WEIGHT_QUANTIZATION, SPARSE_PRUNING, ROW_PRUNING, HEAD_PRUNING, CHANNEL_PRUNING, ACTIVATION_QUANTIZATION]
    methods = sorted(methods, key=lambda x: order.index(list(x[2].keys())[0]))
    for names, related, technique in methods:
        stored = []
        need_mask = any(k in technique for k in (ROW_PRUNING, HEAD_PRUNING, CHANNEL_PRUNING))
        for name in names:
            mask = fix_compression(c_model, name, technique, dim_reduction=need_mask)
            if need_mask:
                stored.append(mask)
        if need_mask and related:
            for i, group in enumerate(related):
                for j, rel_names in enumerate(group):
                    for rn in rel_names:
                        mask = stored[j] if j < len(stored) else None
                        if mask is not None:
                            fix_compression(c_model, rn, technique, mask=mask, dim_reduction=True)
    return model


def student_initialization(student_model, teacher_model, deepspeed_config):
    """Layer reduction: copy the selected teacher layers (and the listed non-layer modules) into the shallower student."""
    cfg = get_compression_config(check_deepspeed_config(deepspeed_config))[LAYER_REDUCTION]
    prefix = cfg[MODULE_NAME_PREFIX]
    teacher_layers = cfg[TEACHER_LAYER]
    student_layers = list(range(len(teacher_layers)))
    others = cfg[OTHER_MODULE_NAME]
    assert len(student_layers) == len(teacher_layers)
    for s, t in zip(student_layers, teacher_layers):
        sm, tm = recursive_getattr(student_model, f"{prefix}.{s}"), recursive_getattr(teacher_model, f"{prefix}.{t}")
        for ps, pt in zip(sm.parameters(), tm.parameters()):
            ps.data.copy_(pt.data)
    for name in others:
        sm, tm = recursive_getattr(student_model, name), recursive_getattr(teacher_model, name)
        for ps, pt in zip(sm.parameters(), tm.parameters()):
            ps.data.copy_(pt.data)
