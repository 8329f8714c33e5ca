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


def _parse_size(size) -> int:
    """``"5GB"`` / ``"200MB"`` / ``"1GiB"`` / int -> bytes."""
    if isinstance(size, int):
        return size
    m = re.fullmatch(r"\s*([0-9.]+)\s*([KMGT]?i?B?)\s*", str(size), re.I)
    if not m:
        raise ValueError(f"cannot parse size {size!r}")
    num, unit = float(m.group(1)), m.group(2).upper()
    base = 1024 if "I" in unit else 1000
    power = {"": 0, "B": 0, "K": 1, "M": 2, "G": 3, "T": 4}[unit.replace("I", "").replace("B", "") or ""]
    return int(num * base ** power)


def _shard_state_dict(sd, max_shard_bytes):
    """Greedy split in insertion order: a new shard starts when the next tensor would exceed ``max_shard_bytes``."""
    shards, cur, cur_bytes = [], OrderedDict(), 0
    for name, t in sd.items():
        nbytes = t.numel() * t.element_size() if torch.is_tensor(t) else 0
        if cur and cur_bytes + nbytes > max_shard_bytes:
            shards.append(cur)
            cur, cur_bytes = OrderedDict(), 0
        cur[name] = t
        cur_bytes += nbytes
    if cur or not shards:
        shards.append(cur)
    return shards


def convert_zero_checkpoint_to_fp32_state_dict(checkpoint_dir, output_dir, max_shard_size="5GB",
                                               safe_serialization=False, tag=None, exclude_frozen_parameters=False):
    """Consolidate a ZeRO-1/2/3 checkpoint into fp32 weights loadable without this framework.

    ``output_dir`` ending in ``.bin`` / ``.pt`` / ``.pth`` / ``.safetensors`` writes that single file (the older
    calling convention).  Otherwise a directory is written the way HuggingFace shards checkpoints:
    ``pytorch_model.bin`` (or ``model.safetensors``) when everything fits ``max_shard_size``, else
    ``pytorch_model-0000i-of-0000n.bin`` plus ``pytorch_model.bin.index.json`` with the ``weight_map``."""
    import json
    sd = get_fp32_state_dict_from_zero_checkpoint(checkpoint_dir, tag, exclude_frozen_parameters)
    save = torch.save
    if safe_serialization:
        try:
            from safetensors.torch import save_file
        except ImportError:
            print("If you want to use `safe_serialization`, please `pip install safetensors`")
            raise
        save = lambda obj, path: save_file({k: v.contiguous() for k, v in obj.items()}, path, metadata={"format": "pt"})  # noqa: E731
    if str(output_dir).endswith((".bin", ".pt", ".pth", ".safetensors")):
        print(f"Saving fp32 state dict to {output_dir}")
        save(sd, output_dir)
        return sd
    os.makedirs(output_dir, exist_ok=True)
    base = "model.safetensors" if safe_serialization else "pytorch_model.bin"
    if safe_serialization:                       # safetensors refuses aliased tensors: materialise ties
        seen_ptr = set()
        for k in list(sd):
            ptr = sd[k].untyped_storage().data_ptr()
            if ptr in seen_ptr:
                sd[k] = sd[k].clone()
            seen_ptr.add(sd[k].untyped_storage().data_ptr())
    shards = _shard_state_dict(sd, _parse_size(max_shard_size))
    if len(shards) == 1:
        print(f"Saving fp32 state dict to {os.path.join(output_dir, base)}")
        save(shards[0], os.path.join(output_dir, base))
        return sd
    stem, ext = base.rsplit(".", 1)
    weight_map, total = OrderedDict(), 0
    for i, shard in enumerate(shards):
        fname = f"{stem}-{i + 1:05d}-of-{len(shards):05d}.{ext}"
        save(shard, os.path.join(output_dir, fname))
        for k, v in shard.items():
            weight_map[k] = fname
            total += v.numel() * v.element_size()
    index = {"metadata": {"total_size": total}, "weight_map": weight_map}
    with open(os.path.join(output_dir, base + ".index.json"), "w", encoding="utf-8") as f:
        f.write(json.dumps(index, indent=2, sort_keys=True) + "\n")
    print(f"Saved {len(shards)} shards and {base}.index.json to {output_dir}")
    return sd


def load_state_dict_from_zero_checkpoint(model, checkpoint_dir, tag=None):
    sd = get_fp32_state_dict_from_zero_checkpoint(checkpoint_dir, tag)
    model = model.cpu()
    model.load_state_dict(sd, strict=False)
    return model


if __name__ == "__main__":
    parser = argparse.ArgumentParser()
    parser.add_argument("checkpoint_dir", type=str, help="path to the checkpoint folder (the one holding `latest`)")
    parser.add_argument("output_dir", type=str,
                        help="directory for the fp32 state_dict files (or a single .bin/.pt/.safetensors file)")
    parser.add_argument("--max_shard_size", type=str, default="5GB",
                        help="maximum size of one output file, e.g. 5GB; larger checkpoints are split and indexed")
    parser.add_argument("--safe_serialization", default=False, action="store_true",
                        help="write safetensors instead of pickled torch files")
    parser.add_argument("-t", "--tag", type=str, default=None, help="checkpoint tag, e.g. global_step14")
    parser.add_argument("--exclude_frozen_parameters", action="store_true", help="exclude frozen parameters")
    parser.add_argument("-d", "--debug", action="store_true", help="enable debug")
    a = parser.parse_args()
    convert_zero_checkpoint_to_fp32_state_dict(a.checkpoint_dir, a.output_dir, max_shard_size=a.max_shard_size,
                                               safe_serialization=a.safe_serialization, tag=a.tag,
                                               exclude_frozen_parameters=a.exclude_frozen_parameters)
