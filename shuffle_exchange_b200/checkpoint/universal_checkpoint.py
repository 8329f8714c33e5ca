"""Loading side of the *universal checkpoint*: one folder per parameter holding FULL (unsharded, un-TP-sliced) fp32
weight and optimizer moments, so a run can resume under any data/tensor/pipeline-parallel layout.

Reference: ``deepspeed/checkpoint/universal_checkpoint.py:13-160`` (``load_hp_checkpoint_state``: read
``<folder>/{fp32,exp_avg,exp_avg_sq}.pt``, slice for the TP rank along ``cat_dim`` — with special handling of padded
vocabularies and multi-sub-parameter tensors —, then narrow to this rank's fragment of the flat partition).
"""
from __future__ import annotations

import os
import re
from typing import Optional

import torch

from .constants import (CAT_DIM, FP32_WEIGHT_KEY, PARAM, PARAM_N_SUB_PARAMS, SUB_PARAM_SHAPE, VOCAB_TENSOR)


def _tp_slice(full: torch.Tensor, ckpt: dict, tp_rank: int, tp_world: int, target_shape) -> torch.Tensor:
    """Cut the tensor-parallel slice of a full tensor according to the metadata stored with it."""
    if tp_world == 1 or tuple(full.shape) == tuple(target_shape):
        return full
    dim = ckpt.get(CAT_DIM, 0)
    n_sub = ckpt.get(PARAM_N_SUB_PARAMS, 1)
    if ckpt.get(VOCAB_TENSOR, False):
        # vocabulary tensors were saved unpadded; re-pad to a multiple of tp_world before slicing
        rows = target_shape[0] * tp_world
        if full.shape[0] < rows:
            pad = torch.zeros((rows - full.shape[0],) + tuple(full.shape[1:]), dtype=full.dtype)
            full = torch.cat([full, pad], 0)
        return full.chunk(tp_world, dim=0)[tp_rank]
    if n_sub > 1:
        # e.g. fused gate|up or q|k|v: slice every sub-parameter separately, then re-fuse
        subs = full.chunk(n_sub, dim=dim)
        return torch.cat([s.chunk(tp_world, dim=dim)[tp_rank] for s in subs], dim=dim)
    sub_shape = ckpt.get(SUB_PARAM_SHAPE)
    if sub_shape is not None:
        sizes = sub_shape.sizes if hasattr(sub_shape, "sizes") else sub_shape
        subs = torch.split(full, list(sizes), dim=dim)
        return torch.cat([s.chunk(tp_world, dim=dim)[tp_rank] for s in subs], dim=dim)
    return full.chunk(tp_world, dim=dim)[tp_rank]


def load_hp_checkpoint_state(optimizer, name: str, lp_param: torch.Tensor, folder: str, tp_rank: int = 0,
                             tp_world_size: int = 1) -> Optional[int]:
    """Fill this rank's shard of ``name`` (fp32 master + every optimizer state found in ``folder``) from full tensors.
    ``optimizer`` is one of the framework's sharded optimizers; the parameter is located through its flat group.
    Returns the optimizer step stored with the parameter (or None)."""
    link = getattr(lp_param, "_sxb_link", None)
    assert link is not None, f"{name} is not managed by a sharded optimizer"
    fg, idx = link.group, link.index
    off, numel, shape = fg.offsets[idx], fg.numels[idx], fg.shapes[idx]
    lo, hi = fg.rank * fg.part, (fg.rank + 1) * fg.part
    a, b = max(off, lo), min(off + numel, hi)           # overlap of the parameter with my partition
    step = None
    files = [f for f in os.listdir(folder) if f.endswith(".pt")]
    for fname in files:
        key = fname[:-3]
        ckpt = torch.load(os.path.join(folder, fname), map_location="cpu", weights_only=False)
        if key == "step":
            step = ckpt if not isinstance(ckpt, dict) else ckpt.get(PARAM, ckpt)
            continue
        full = ckpt[PARAM] if isinstance(ckpt, dict) else ckpt
        if not torch.is_tensor(full):
            continue
        full = _tp_slice(full, ckpt if isinstance(ckpt, dict) else {}, tp_rank, tp_world_size, shape)
        assert full.numel() == numel, f"{name}/{key}: checkpoint has {full.numel()} elements, parameter has {numel}"
        if b <= a:
            continue
        piece = full.reshape(-1)[a - off: b - off].float()
        if key == FP32_WEIGHT_KEY:
            fg.hp_param.data[a - lo: b - lo].copy_(piece)
        else:
            # optimizer state is created lazily at the first step: materialise a zero shard for every saved moment
            st = optimizer.optimizer.state[fg.hp_param]
            if key not in st or not torch.is_tensor(st[key]) or st[key].numel() != fg.part:
                st[key] = torch.zeros_like(fg.hp_param.data)
                st.setdefault("step", 0)
            st[key][a - lo: b - lo].copy_(piece.to(st[key].dtype))
    return step


def enable_universal_checkpoint(param_list):
    """Mark parameters so that engines know they can be restored from a universal checkpoint (kept for API parity:
    in this framework every sharded optimizer supports it through :func:`load_hp_checkpoint_state`)."""
    for p in param_list:
        p.load_hp_checkpoint_state = load_hp_checkpoint_state


def load_universal_into_engine(engine, load_dir: str, tag: str, load_optimizer_states: bool = True,
                               load_lr_scheduler_states: bool = True):
    """Engine-side entry (``checkpoint.load_universal = true``): restore counters / scheduler from the copied
    model-state file, then every parameter's fp32 master (and moments) from ``<tag>/zero/<name>/``, re-deriving the
    low-precision shards and the gathered parameters from the masters."""
    import glob

    from ..utils.logging import logger
    folder = os.path.join(load_dir, str(tag))
    zero_dir = os.path.join(folder, "zero")
    if not os.path.isdir(zero_dir):
        raise ValueError(f"{folder} is not a universal checkpoint (no zero/ folder)")
    model_files = sorted(glob.glob(os.path.join(folder, "*_model_states.pt")))
    ckpt = torch.load(model_files[0], map_location="cpu", weights_only=False) if model_files else {}
    engine.global_steps = ckpt.get("global_steps", 0)
    engine.global_samples = ckpt.get("global_samples", engine.global_steps * engine.train_batch_size())
    engine.skipped_steps = ckpt.get("skipped_steps", 0)
    if load_lr_scheduler_states and engine.lr_scheduler is not None and ckpt.get("lr_scheduler") is not None:
        engine.lr_scheduler.load_state_dict(ckpt["lr_scheduler"])
    opt = engine.optimizer
    assert hasattr(opt, "groups"), "universal checkpoints restore sharded (ZeRO / bf16) optimizers"
    tp_rank = tp_world = None
    mpu = getattr(engine, "mpu", None)
    if mpu is not None:
        from ..utils.bwc import bwc_tensor_model_parallel_rank, bwc_tensor_model_parallel_world_size
        tp_rank, tp_world = bwc_tensor_model_parallel_rank(mpu), bwc_tensor_model_parallel_world_size(mpu)
    steps = []
    for fg in opt.groups:
        for name, p in zip(fg.names, fg.params):
            pdir = os.path.join(zero_dir, name)
            if not os.path.isdir(pdir):
                logger.warning(f"universal checkpoint has no entry for {name}")
                continue
            if not load_optimizer_states:
                # weights only: hide the moment files from the loader
                ck = torch.load(os.path.join(pdir, f"{FP32_WEIGHT_KEY}.pt"), map_location="cpu", weights_only=False)
                tmp = {FP32_WEIGHT_KEY: ck}
            st = load_hp_checkpoint_state(opt, name, p, pdir, tp_rank or 0, tp_world or 1)
            if st is not None:
                steps.append(int(st))
        with torch.no_grad():
            fg.lp_shard.copy_(fg.hp_param.data.to(fg.lp_shard.device))
        if load_optimizer_states and steps:
            state = opt._optimizer_state_for(fg)
            if "step" in state:
                state["step"] = (torch.tensor(float(max(steps))) if torch.is_tensor(state["step"]) else max(steps))
    # make the module see the restored weights
    if hasattr(opt, "_refresh_gathered"):
        opt._refresh_gathered()
    else:
        for fg in opt.groups:
            if fg.params_available():
                fg.refresh_full_from_shard()
    skip = {"module", "buffer_names", "optimizer", "param_shapes", "frozen_param_shapes", "shared_params",
            "frozen_param_fragments", "lr_scheduler", "data_sampler", "random_ltd", "sparse_tensor_module_names",
            "skipped_steps", "global_steps", "global_samples", "dp_world_size", "mp_world_size", "ds_config",
            "ds_version", "num_experts"}
    return folder, {k: v for k, v in ckpt.items() if k not in skip}
