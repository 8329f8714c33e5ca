"""ZeRO-3 leaf modules (reference: deepspeed/utils/z3_leaf_module.py:10-93): a flagged module is fetched /
released as one unit, its children get no hooks of their own (needed for MoE blocks whose active
experts vary per step).  In this framework the flag simply makes the module a ZeRO-3 *unit*."""
from typing import List, Type, Union

import torch


def z3_leaf_module(model: torch.nn.Module) -> bool:
    return getattr(model, "_z3_leaf", False)


def z3_leaf_parameter(model: torch.nn.Parameter) -> bool:
    return getattr(model, "ds_z3_leaf_module", None) is not None


def get_z3_leaf_modules(model: torch.nn.Module) -> List[torch.nn.Module]:
    return [m for m in model.modules() if z3_leaf_module(m)]


def set_z3_leaf_module(model: torch.nn.Module, flag: bool):
    """Flag (or unflag) this one module; its parameters remember which unit they belong to."""
    model._z3_leaf = flag
    for p in model.parameters():
        p.ds_z3_leaf_module = model if flag else None


def _do_set_z3_leaf_modules(model, leaf_module_classes, flag: bool):
    assert all(isinstance(c, (type, str)) for c in leaf_module_classes), \
        f"leaf_module_classes must be a list of types or names, got {leaf_module_classes}"
    leaf_modules = []

    def _set(m: torch.nn.Module):
        for c in leaf_module_classes:
            if (isinstance(c, type) and m.__class__ == c) or (isinstance(c, str) and m.__class__.__name__ == c):
                set_z3_leaf_module(m, flag)
                leaf_modules.append(m)

    model.apply(_set)
    if len(leaf_modules) == 0:
        raise ValueError(f"No modules of type {leaf_module_classes} found in model {model}")
    return leaf_modules


def set_z3_leaf_modules(model: torch.nn.Module, leaf_module_classes: Union[List[Type], List[str]]):
    return _do_set_z3_leaf_modules(model, leaf_module_classes, True)


def unset_z3_leaf_modules(model: torch.nn.Module, leaf_module_classes: List[Type]):
    return _do_set_z3_leaf_modules(model, leaf_module_classes, False)
