#!/usr/bin/env python
"""Convert a ZeRO checkpoint into the universal format.

Reference: ``deepspeed/checkpoint/ds_to_universal.py:60-560`` (extract per-rank fragments -> merge TP slices ->
``<out>/zero/<param>/{fp32,exp_avg,exp_avg_sq}.pt`` + ``<out>/zero/optimizer_state.pt`` + copied model-state files +
``latest_universal``).  The flat-group checkpoints of this framework store explicit ``param_offsets`` /
``param_shapes`` tables, so extraction is one concatenation of the rank shards per state followed by slicing —
no temporary fragment files.

    python -m shuffle_exchange_b200.checkpoint.ds_to_universal --input_folder ckpt/global_step10 --output_folder ckpt/global_step10_universal
"""
from __future__ import annotations

import argparse
import glob
import os
import re
import shutil

import torch

from .constants import (BASE_OPTIMIZER_STATE, CAT_DIM, FP32_FLAT_GROUPS, FP32_WEIGHT_KEY, OPTIMIZER_STATE_DICT, PARAM,
                        SINGLE_PARTITION_OF_FP32_GROUPS, UNIVERSAL_CHECKPOINT_INFO, UNIVERSAL_CHECKPOINT_VERSION_KEY,
                        UNIVERSAL_CHECKPOINT_VERSION_VALUE, ZERO_STAGE)


def _natural(s):
    return [int(t) if t.isdigit() else t for t in re.split(r"(\d+)", s)]


def parse_arguments(argv=None):
    p = argparse.ArgumentParser()
    p.add_argument("--input_folder", type=str, required=True, help="ZeRO checkpoint folder (the tag directory)")
    p.add_argument("--output_folder", type=str, required=True, help="output universal checkpoint folder")
    p.add_argument("--num_extract_workers", default=4, type=int)
    p.add_argument("--num_merge_workers", default=2, type=int)
    p.add_argument("--keep_temp_folder", action="store_true")
    p.add_argument("--no_strict", dest="strict", action="store_false")
    p.add_argument("--inject_missing_state", action="store_true")
    return p.parse_args(argv)


def _rank_states(folder):
    files = sorted(glob.glob(os.path.join(folder, "*_optim_states.pt")), key=_natural)
    if not files:
        raise FileNotFoundError(f"no *_optim_states.pt under {folder}")
    return [torch.load(f, map_location="cpu", weights_only=False)[OPTIMIZER_STATE_DICT] for f in files]


def _group_states(sd, gi):
    """(fp32 shard, {state name: shard}, step) of group ``gi`` in one rank's optimizer state dict."""
    if sd[ZERO_STAGE] == 3:
        st = sd[OPTIMIZER_STATE_DICT]["state"][gi]
        fp32 = sd[FP32_FLAT_GROUPS][gi]
    else:
        st = sd[BASE_OPTIMIZER_STATE][gi]
        fp32 = sd[SINGLE_PARTITION_OF_FP32_GROUPS][gi]
    tensors = {k: v for k, v in st.items() if torch.is_tensor(v) and v.numel() > 1}
    return fp32, tensors, st.get("step", 0)


def convert(input_folder: str, output_folder: str, strict: bool = True) -> None:
    states = _rank_states(input_folder)
    pc = states[0]["partition_count"]
    n_groups = len(states[0]["param_offsets"])
    zero_dir = os.path.join(output_folder, "zero")
    os.makedirs(zero_dir, exist_ok=True)
    for gi in range(n_groups):
        world = pc[gi] if isinstance(pc, (list, tuple)) else pc
        ranks = states[:world]                       # replicas (slices / MiCS) repeat the same shard files
        per_rank = [_group_states(sd, gi) for sd in ranks]
        flat = {FP32_WEIGHT_KEY: torch.cat([r[0].reshape(-1).float() for r in per_rank])}
        for key in per_rank[0][1]:
            flat[key] = torch.cat([r[1][key].reshape(-1).float() for r in per_rank])
        step = per_rank[0][2]
        shapes = states[0]["param_shapes"][gi]
        for name, (off, numel) in states[0]["param_offsets"][gi].items():
            folder = os.path.join(zero_dir, name)
            os.makedirs(folder, exist_ok=True)
            for key, buf in flat.items():
                t = buf[off: off + numel].view(tuple(shapes[name])).clone()
                torch.save({PARAM: t, CAT_DIM: 0}, os.path.join(folder, f"{key}.pt"))
            torch.save(int(step) if not torch.is_tensor(step) else int(step.item()), os.path.join(folder, "step.pt"))
    # global (non-tensor) optimizer state of rank 0
    glob_state = {k: v for k, v in states[0].items()
                  if k not in (FP32_FLAT_GROUPS, SINGLE_PARTITION_OF_FP32_GROUPS, BASE_OPTIMIZER_STATE, OPTIMIZER_STATE_DICT)}
    glob_state[UNIVERSAL_CHECKPOINT_INFO] = {UNIVERSAL_CHECKPOINT_VERSION_KEY: UNIVERSAL_CHECKPOINT_VERSION_VALUE}
    torch.save(glob_state, os.path.join(zero_dir, "optimizer_state.pt"))
    for f in glob.glob(os.path.join(input_folder, "*_model_states.pt")):
        shutil.copy2(f, output_folder)
    parent, tag = os.path.split(os.path.normpath(output_folder))
    with open(os.path.join(parent, "latest_universal"), "w") as fh:
        fh.write(tag)


def main(args=None):
    a = args if args is not None and not isinstance(args, list) else parse_arguments(args)
    print(f"Convert DeepSpeed Checkpoint to Universal Checkpoint\n  {a.input_folder} -> {a.output_folder}")
    convert(a.input_folder, a.output_folder, strict=a.strict)
    print("*** Done!")


if __name__ == "__main__":
    main()
