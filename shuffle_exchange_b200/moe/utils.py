"""MoE parameter bookkeeping (reference: deepspeed/moe/utils.py:12-182)."""
from collections import defaultdict
from typing import Dict, List, Tuple

import torch

from .layer import MoE


def has_moe_layers(m: torch.nn.Module) -> Tuple[bool, int]:
    found, num_experts = False, 0
    for module in m.modules():
        if isinstance(module, MoE):
            found = True
            num_experts = module.num_experts
            break
    return found, num_experts


def is_moe_param(param: torch.Tensor) -> bool:
    return hasattr(param, "allreduce") and not param.allreduce


def split_params_into_shared_and_expert_params(params):
    shared, expert = [], []
    for p in params:
        (expert if is_moe_param(p) else shared).append(p)
    return shared, expert


def split_params_grads_into_shared_and_expert_params(group):
    shared, expert = [], []
    for p in group:
        if p.grad is not None:
            (expert if is_moe_param(p) else shared).append(p.grad.to(p.dtype))
    return shared, expert


def split_params_into_different_moe_groups_for_optimizer(param_groups, max_group_size=178956971):
    """Move expert parameters of every group into new groups named after their expert-parallel group
    (``moe: True``, ``name: ep_size_N``) so ZeRO shards them over the matching expert-DP group."""
    if isinstance(param_groups, tuple):
        param_groups = list(param_groups)
    elif isinstance(param_groups, dict):
        param_groups = [param_groups]
    elif not isinstance(param_groups, list):
        raise ValueError(f"Unknown param group type of {type(param_groups)}")
    data_parallel_group_names = set()
    for pg in param_groups:
        for p in pg["params"]:
            if is_moe_param(p):
                data_parallel_group_names.add(p.group_name)
    names = sorted(data_parallel_group_names)
    moe_groups: Dict[str, Dict[str, dict]] = {n: {} for n in names}
    for pg in param_groups:
        for n in names:
            g = {k: v for k, v in pg.items() if k not in ("name", "params")}
            g["moe"] = True
            g["name"] = n
            g["params"] = []
            moe_groups[n][pg.get("name", id(pg))] = g
    for pg in param_groups:
        keep = []
        for p in pg["params"]:
            if is_moe_param(p):
                moe_groups[p.group_name][pg.get("name", id(pg))]["params"].append(p)
            else:
                keep.append(p)
        pg["params"] = keep
    for n in names:
        for g in moe_groups[n].values():
            if not g["params"]:
                continue
            if max_group_size is None:
                param_groups.append(g)
                continue
            cur, size = [], 0
            chunks = []
            for p in g["params"]:
                if size + p.numel() > max_group_size and cur:
                    chunks.append(cur)
                    cur, size = [], 0
                cur.append(p)
                size += p.numel()
            if cur:
                chunks.append(cur)
            for c in chunks:
                ng = {k: v for k, v in g.items() if k != "params"}
                ng["params"] = c
                param_groups.append(ng)
    return param_groups


def configure_moe_param_groups(model_parameters):
    assert isinstance(model_parameters, list), "model_parameters must be a list (of tensors or of param-group dicts)"
    if not model_parameters:
        return model_parameters
    if isinstance(model_parameters[0], dict):
        if not any("moe" in g for g in model_parameters):
            return split_params_into_different_moe_groups_for_optimizer(model_parameters)
        return model_parameters
    return split_params_into_different_moe_groups_for_optimizer({"params": model_parameters})


def is_moe_param_group(param_group):
    return param_group.get("moe", False)
