#!/usr/bin/env python
"""Offline consolidation of a ZeRO checkpoint into one fp32 ``state_dict``.

Same role and CLI as the reference script that every checkpoint directory receives a copy of
(deepspeed/utils/zero_to_fp32.py:533-760): ``python zero_to_fp32.py <checkpoint_dir> <output_file> [-t tag]``.
It needs neither the engine nor a GPU: it reads every ``*_optim_states.pt`` file, concatenates the per-rank
fp32 partitions of each flat group and cuts the parameters out using the ``param_offsets`` /
``param_shapes`` tables stored with them (this framework aligns every parameter to 64 elements inside
a group, so offsets are stored explicitly instead of being re-derived as the reference does at :304-316
and :437-488).
"""
import argparse
import glob
import os
import re
from collections import OrderedDict

import torch


def _natural(s):
    return [int(t) if t.isdigit() else t for t in re.split(r"(\d+)", s)]


def get_optim_files(checkpoint_dir):
    files = sorted(glob.glob(os.path.join(checkpoint_dir, "*_optim_states.pt")), key=_natural)
    if not files:
        raise FileNotFoundError(f"can't find *_optim_states.pt files in directory '{checkpoint_dir}'")
    return files


def get_model_state_files(checkpoint_dir):
    files = sorted(glob.glob(os.path.join(checkpoint_dir, "*_model_states.pt")), key=_natural)
    if not files:
        raise FileNotFoundError(f"can't find *_model_states.pt files in directory '{checkpoint_dir}'")
    return files


def _load(path):
    return torch.load(path, map_location="cpu", weights_only=False)


def get_fp32_state_dict_from_zero_checkpoint(checkpoint_dir, tag=None, exclude_frozen_parameters=False):
    """Returns an ``OrderedDict`` name -> fp32 tensor assembled from all ranks' partitions."""
    if tag is None:
        latest = os.path.join(checkpoint_dir, "latest")
        if os.path.isfile(latest):
            with open(latest) as f:
                tag = f.read().strip()
        else:
            raise ValueError(f"Unable to find 'latest' file at {latest}")
    ds_dir = os.path.join(checkpoint_dir, tag)
    if not os.path.isdir(ds_dir):
        raise FileNotFoundError(f"Directory '{ds_dir}' doesn't exist")
    optim_files = get_optim_files(ds_dir)
    states = [_load(f)["optimizer_state_dict"] for f in optim_files]
    stage = states[0]["zero_stage"]
    pc = states[0]["partition_count"]
    world = pc[0] if isinstance(pc, (list, tuple)) else pc
    # with replica slices several replicas write the same partition index: keep the first `world`
    states = states[:world]
    key = "fp32_flat_groups" if stage == 3 else "single_partition_of_fp32_groups"
    n_groups = len(states[0][key])
    out = OrderedDict()
    for gi in range(n_groups):
        flat = torch.cat([s[key][gi].reshape(-1).float() for s in states])
        offsets = states[0]["param_offsets"][gi]
        shapes = states[0]["param_shapes"][gi]
        for name, (off, numel) in offsets.items():
            out[name] = flat[off: off + numel].view(tuple(shapes[name])).clone()
    model_sd = _load(get_model_state_files(ds_dir)[0])
    if not exclude_frozen_parameters:
        frozen = model_sd.get("frozen_param_shapes") or {}
        module = model_sd.get("module") or {}
        for name in frozen:
            if name in module and torch.is_tensor(module[name]) and module[name].numel() > 0:
                out[name] = module[name].float()
    for name in model_sd.get("buffer_names") or []:
        module = model_sd.get("module") or {}
        if name in module and torch.is_tensor(module[name]) and module[name].numel() > 0:
            out[name] = module[name].float()
    for dup, src in (model_sd.get("shared_params") or {}).items():
        if src in out:
            out[dup] = out[src]
    return out


def convert_zero_checkpoint_to_fp32_state_dict(checkpoint_dir, output_file, tag=None, exclude_frozen_parameters=False):
    sd = get_fp32_state_dict_from_zero_checkpoint(checkpoint_dir, tag, exclude_frozen_parameters)
    print(f"Saving fp32 state dict to {output_file}")
    torch.save(sd, output_file)
    return sd


def load_state_dict_from_zero_checkpoint(model, checkpoint_dir, tag=None):
    sd = get_fp32_state_dict_from_zero_checkpoint(checkpoint_dir, tag)
    model = model.cpu()
    model.load_state_dict(sd, strict=False)
    return model


if __name__ == "__main__":
    parser = argparse.ArgumentParser()
    parser.add_argument("checkpoint_dir", type=str, help="path to the checkpoint folder (the one holding `latest`)")
    parser.add_argument("output_file", type=str, help="path to the pytorch fp32 state_dict output file")
    parser.add_argument("-t", "--tag", type=str, default=None, help="checkpoint tag, e.g. global_step14")
    parser.add_argument("--exclude_frozen_parameters", action="store_true")
    a = parser.parse_args()
    convert_zero_checkpoint_to_fp32_state_dict(a.checkpoint_dir, a.output_file, tag=a.tag,
                                               exclude_frozen_parameters=a.exclude_frozen_parameters)
