"""Engine checkpoint save / load (bound as methods of :class:`DeepSpeedEngine`).

On-disk layout and key names follow the reference (deepspeed/runtime/engine.py:2826-3959; table in
SURVEY.md §5.4):

    <save_dir>/latest                                             text file with the newest tag
    <save_dir>/<tag>/mp_rank_{MP:02d}_model_states.pt              module + bookkeeping
    <save_dir>/<tag>/zero_pp_rank_{DP}_mp_rank_{MP:02d}_model_states.pt   (ZeRO-3: every rank)
    <save_dir>/<tag>/[bf16_]zero_pp_rank_{DP}_mp_rank_{MP:02d}_optim_states.pt
    <save_dir>/<tag>/layer_{L}_expert_{E}_mp_rank_{MP:02d}_model_states.pt      (MoE experts)
    <save_dir>/<tag>/expp_rank_{EP}_mp_rank_{MP:02d}_optim_states.pt            (MoE, non-ZeRO)
    <save_dir>/<tag>/zero_to_fp32.py                               offline consolidation script

One conscious fix (SURVEY.md §2.2-A "side effect"): with replica slices the reference names optimizer
files by the rank *inside the slice*, so different replicas overwrite each other.  Here the file index
is the rank in the full sequence-data-parallel group, and a ``slice_rank`` key is stored inside.
"""
from __future__ import annotations

import glob
import hashlib
import os
import shutil
from collections import OrderedDict

import torch

from .. import comm as dist
from ..utils import groups
from ..utils.logging import log_dist, logger
from ..version import __version__

CKPT_LATEST = "latest"
CKPT_LATEST_UNIVERSAL = "latest_universal"


def _mp_rank(engine):
    return 0 if engine.mpu is None else groups.get_model_parallel_rank()


def _get_ckpt_name(self, checkpoints_path, tag, mp_placeholder=None):
    mp = mp_placeholder if mp_placeholder is not None else f"{_mp_rank(self):02d}"
    if self.zero_optimization_partition_weights():
        rank = dist.get_rank(self.optimizer.dp_process_group) if dist.is_initialized() else 0
        name = f"zero_pp_rank_{rank}_mp_rank_{mp}_model_states.pt"
    else:
        name = f"mp_rank_{mp}_model_states.pt"
    return os.path.join(checkpoints_path, str(tag), name)


def _get_rank_zero_ckpt_name(self, checkpoints_path, tag, mp_rank, dp_rank, bf16_mode):
    prefix = "bf16_" if bf16_mode else ""
    return os.path.join(checkpoints_path, str(tag), f"{prefix}zero_pp_rank_{dp_rank}_mp_rank_{mp_rank:02d}_optim_states.pt")


def _get_zero_ckpt_prefix(self, dp_rank, bf16_mode):
    return f'{"bf16_" if bf16_mode else ""}zero_pp_rank_{dp_rank}'


def _zero_file_rank(self):
    """Index used in optimizer-state file names: rank in the full (sequence x data) parallel group."""
    return groups._get_sequence_data_parallel_rank()


def _get_zero_ckpt_name(self, checkpoints_path, tag):
    bf16_mode = self.bfloat16_enabled()
    return _get_rank_zero_ckpt_name(self, checkpoints_path, tag, _mp_rank(self), _zero_file_rank(self), bf16_mode)


def _get_optimizer_ckpt_name(self, checkpoints_path, tag, expp_rank):
    return os.path.join(checkpoints_path, str(tag), f"expp_rank_{expp_rank}_mp_rank_{_mp_rank(self):02d}_optim_states.pt")


def _get_expert_ckpt_name(checkpoints_path, layer_id, expert_id, tag, mpu=None):
    mp_rank = 0 if mpu is None else groups.get_model_parallel_rank()
    if layer_id <= -1:
        name = f"expert_{expert_id}_mp_rank_{mp_rank:02d}_model_states.pt"   # legacy
    else:
        name = f"layer_{layer_id}_expert_{expert_id}_mp_rank_{mp_rank:02d}_model_states.pt"
    return os.path.join(checkpoints_path, "" if tag is None else str(tag), name)


def _checkpoint_tag_validation(self, tag):
    if not self.checkpoint_tag_validation_enabled() or not dist.is_initialized() or self.world_size == 1:
        return
    digest = hashlib.sha1(str(tag).encode()).digest()
    t = torch.tensor(list(digest), dtype=torch.int64, device=self.device)
    mx, mn = t.clone(), t.clone()
    dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    dist.all_reduce(mn, op=dist.ReduceOp.MIN)
    valid = bool(torch.equal(mx, t) and torch.equal(mn, t))
    msg = (f"[rank={dist.get_rank()}] The checkpoint tag name '{tag}' is not consistent across all ranks. Including "
           "rank unique information in checkpoint tag could cause issues when restoring with different world sizes.")
    if self.checkpoint_tag_validation_fail():
        assert valid, msg
    elif not valid:
        logger.warning(msg)


def _param_shapes(self):
    """List (one per flat group) of name -> shape in flat order (reference key ``param_shapes``)."""
    opt = self.optimizer
    if hasattr(opt, "groups"):
        return [OrderedDict(zip(fg.names, [torch.Size(s) for s in fg.shapes])) for fg in opt.groups]
    return []


def save_checkpoint(self, save_dir, tag=None, client_state={}, save_latest=True, exclude_frozen_parameters=False):
    """All ranks must call.  Returns True."""
    if tag is None:
        tag = f"global_step{self.global_steps}"
    tag = str(tag)
    if self.checkpoint_engine.is_decoupled():
        self._commit_decoupled_checkpoint()          # two saves with no optimizer step in between
    self.checkpoint_engine.create(tag)
    self._checkpoint_tag_validation(tag)
    os.makedirs(os.path.join(save_dir, tag), exist_ok=True)
    self._curr_ckpt_path = os.path.join(save_dir, tag)
    if dist.is_initialized():
        dist.barrier()
    if self.has_moe_layers:
        self._save_moe_checkpoint(save_dir, tag, client_state=client_state,
                                  exclude_frozen_parameters=exclude_frozen_parameters)
    elif self.save_non_zero_checkpoint or self.zero_optimization_partition_weights():
        _save_model_states(self, save_dir, tag, client_state, exclude_frozen_parameters)
    if getattr(self, "save_zero_checkpoint", False) and hasattr(self.optimizer, "state_dict") and hasattr(self.optimizer, "groups"):
        self._save_zero_checkpoint(save_dir, tag)
    if self.zero_has_nvme_offload():
        from .swap_tensor.utils import copy_swap_folder
        copy_swap_folder(self, os.path.join(save_dir, tag, "offloaded_tensors"))
    if not self.checkpoint_engine.is_decoupled():
        self.checkpoint_engine.commit(tag)
        if save_latest and self.global_rank == 0:
            with open(os.path.join(save_dir, CKPT_LATEST), "w") as fd:
                fd.write(tag)
    else:
        from .checkpoint_engine.checkpoint_engine import CheckpointCommitInfo
        self.checkpoint_engine.set_commit_info(CheckpointCommitInfo(tag=tag, save_dir=save_dir, save_latest=save_latest))
    if dist.is_initialized():
        dist.barrier()
    return True


def _commit_decoupled_checkpoint(self):
    """Finish the checkpoint a decoupled engine has been writing in the background: wait until it is durable, then move
    ``latest``.  Called at the next optimizer step and from ``destroy`` (reference ``engine.py:3437-3452``)."""
    info = self.checkpoint_engine.get_commit_info()
    if info is None:
        return
    self.checkpoint_engine.commit(info)
    if info.save_latest and self.global_rank == 0:
        with open(os.path.join(info.save_dir, CKPT_LATEST), "w") as fd:
            fd.write(info.tag)
    if dist.is_initialized():
        dist.barrier()


def _model_state_payload(self, client_state, exclude_frozen_parameters, module_sd):
    opt = self.optimizer
    zero_opt = hasattr(opt, "groups")
    frozen = OrderedDict((n, tuple(p.shape)) for n, p in self.module.named_parameters() if not p.requires_grad)
    state = dict(
        module=module_sd,
        buffer_names=[n for n, _ in self.module.named_buffers()],
        optimizer=(None if zero_opt or opt is None else opt.state_dict()),
        param_shapes=_param_shapes(self) if zero_opt else None,
        frozen_param_shapes=frozen if zero_opt and not exclude_frozen_parameters else None,
        shared_params=_shared_params(self.module) if zero_opt else None,
        frozen_param_fragments=None,
        lr_scheduler=self.lr_scheduler.state_dict() if self.lr_scheduler is not None else None,
        data_sampler=(self.training_dataloader.data_sampler.state_dict()
                      if (self.training_dataloader is not None and self.curriculum_learning_enabled()) else None),
        random_ltd=None,
        sparse_tensor_module_names=[],
        skipped_steps=self.skipped_steps,
        global_steps=self.global_steps,
        global_samples=self.global_samples,
        dp_world_size=self.seq_dp_world_size,
        mp_world_size=self.mp_world_size,
        ds_config=self._config._param_dict,
        ds_version=__version__,
    )
    state.update(client_state)
    return state


def _shared_params(module):
    seen, shared = {}, {}
    for n, p in module.named_parameters(remove_duplicate=False):
        if id(p) in seen:
            shared[n] = seen[id(p)]
        else:
            seen[id(p)] = n
    return shared


def _save_model_states(self, save_dir, tag, client_state, exclude_frozen_parameters):
    path = self._get_ckpt_name(save_dir, tag)
    if self.zero_optimization_partition_weights():
        # every rank writes its own file; module weights are placeholders (the fp32 shards in the
        # optim file are authoritative) unless 16-bit gathering on save is requested
        if self.zero_gather_16bit_weights_on_model_save():
            module_sd = self.module_state_dict(exclude_frozen_parameters=exclude_frozen_parameters)
        else:
            module_sd = OrderedDict((k, torch.empty(0, dtype=v.dtype)) for k, v in self.module.state_dict().items())
    else:
        if not self.save_non_zero_checkpoint:
            return
        module_sd = self.module_state_dict(exclude_frozen_parameters=exclude_frozen_parameters)
    state = _model_state_payload(self, client_state, exclude_frozen_parameters, module_sd)
    log_dist(f"Saving model checkpoint: {path}", ranks=[0, 1])
    self.checkpoint_engine.save(state, path)


def _copy_recovery_script(self, save_path):
    src = os.path.join(os.path.dirname(os.path.dirname(__file__)), "utils", "zero_to_fp32.py")
    dst = os.path.join(save_path, "zero_to_fp32.py")
    try:
        shutil.copyfile(src, dst)
        os.chmod(dst, os.stat(dst).st_mode | 0o111)
    except Exception as e:  # pragma: no cover
        logger.warning(f"could not copy zero_to_fp32.py: {e}")


def _save_zero_checkpoint(self, save_path, tag):
    path = self._get_zero_ckpt_name(save_path, tag)
    sd = dict(optimizer_state_dict=self.optimizer.state_dict(), ds_config=self._config._param_dict,
              ds_version=__version__)
    sd["slice_rank"] = dist.get_rank(self.optimizer.dp_process_group) if dist.is_initialized() else 0
    self.checkpoint_engine.save(sd, path)
    if self.global_rank == 0:
        self._copy_recovery_script(os.path.join(save_path, str(tag)))
    log_dist(f"zero checkpoint saved {path}", ranks=[0])


def _save_moe_checkpoint(self, save_dir, tag, client_state={}, exclude_frozen_parameters=False):
    from ..moe.layer import MoE
    from ..moe.utils import has_moe_layers
    moe_layer_id = 0
    full_sd = self.module_state_dict(exclude_frozen_parameters=exclude_frozen_parameters)
    expert_keys = set()
    for name, module in self.module.named_modules():
        if not isinstance(module, MoE):
            continue
        group_name = module.expert_group_name
        num_local = module.num_local_experts
        expp_rank = groups._get_expert_parallel_rank(group_name)
        exp_dp_rank = groups._get_expert_data_parallel_rank(group_name)
        prefix = f"{name}.deepspeed_moe.experts.deepspeed_experts."
        local_sd = {k: v for k, v in full_sd.items() if k.startswith(prefix)}
        expert_keys.update(local_sd.keys())
        if exp_dp_rank == 0:
            for local_id in range(num_local):
                gid = expp_rank * num_local + local_id
                lp = f"{prefix}{local_id}."
                esd = {k.replace(lp, f"{prefix}{gid}."): v.clone().detach() for k, v in local_sd.items()
                       if k.startswith(lp)}
                self.checkpoint_engine.save(esd, _get_expert_ckpt_name(save_dir, moe_layer_id, gid, tag, self.mpu))
        moe_layer_id += 1
    largest = groups._get_max_expert_size_name()
    expp_rank = groups._get_expert_parallel_rank(largest)
    exp_dp_rank = groups._get_expert_data_parallel_rank(largest)
    if exp_dp_rank == 0 and not hasattr(self.optimizer, "groups"):
        osd = {"optimizer": self.optimizer.state_dict() if self.optimizer and not self.zero_optimization() else None}
        self.checkpoint_engine.save(osd, self._get_optimizer_ckpt_name(save_dir, tag, expp_rank))
    # non-expert state: one writer per model-parallel rank — the first rank of the largest expert group's
    # (expert-parallel, expert-data-parallel) grid; its file is mp_rank_XX (reference: engine.py:3523-3545)
    if expp_rank == 0 and exp_dp_rank == 0:
        non_expert = OrderedDict((k, v) for k, v in full_sd.items() if k not in expert_keys)
        state = _model_state_payload(self, client_state, exclude_frozen_parameters, non_expert)
        state["num_experts"] = self.num_experts
        self.checkpoint_engine.save(state, self._get_ckpt_name(save_dir, tag))


# ------------------------------------------------------------------------------------ loading
def load_checkpoint(self, load_dir, tag=None, load_module_strict=True, load_optimizer_states=True,
                    load_lr_scheduler_states=True, load_module_only=False, custom_load_fn=None):
    """Returns ``(load_path, client_state)`` or ``(None, None)`` when nothing could be loaded."""
    if tag is None:
        latest = CKPT_LATEST_UNIVERSAL if self.load_universal_checkpoint() else CKPT_LATEST
        latest_path = os.path.join(load_dir, latest)
        if os.path.isfile(latest_path):
            with open(latest_path, "r") as fd:
                tag = fd.read().strip()
        else:
            if self.load_universal_checkpoint():
                raise ValueError(f"Invalid for universal checkpoint: {latest_path} does not exist")
            logger.warning(f"Unable to find latest file at {latest_path}, if trying to load latest checkpoint please "
                           "ensure this file exists or pass an explicit checkpoint tag when loading a checkpoint.")
            return None, None
    if self.load_universal_checkpoint():
        from ..checkpoint.universal_checkpoint import load_universal_into_engine
        return load_universal_into_engine(self, load_dir, tag, load_optimizer_states, load_lr_scheduler_states)
    path = self._get_ckpt_name(load_dir, tag)
    self._curr_ckpt_path = os.path.join(load_dir, str(tag))
    if not os.path.exists(path):
        alt = os.path.join(load_dir, str(tag), f"mp_rank_{_mp_rank(self):02d}_model_states.pt")
        if os.path.exists(alt):
            path = alt
        else:
            logger.warning(f"Client provided checkpoint load path: {path} does not exist")
            return None, None
    ckpt = self.checkpoint_engine.load(path, map_location="cpu")
    is_z3 = self.zero_optimization_partition_weights()
    if self.has_moe_layers:
        _load_moe_experts(self, load_dir, tag, ckpt)
    if not is_z3 or (ckpt.get("module") and any(v.numel() > 0 for v in ckpt["module"].values() if torch.is_tensor(v))):
        self.load_module_state_dict(ckpt, strict=load_module_strict, custom_load_fn=custom_load_fn)
    self.loaded_checkpoint_dp_world_size = ckpt.get("dp_world_size")
    self.loaded_checkpoint_mp_world_size = ckpt.get("mp_world_size")
    if not load_module_only:
        self.global_steps = ckpt.get("global_steps", 0)
        self.global_samples = ckpt.get("global_samples", self.global_steps * self.train_batch_size())
        self.skipped_steps = ckpt.get("skipped_steps", 0)
        if load_lr_scheduler_states and self.lr_scheduler is not None and ckpt.get("lr_scheduler") is not None:
            self.lr_scheduler.load_state_dict(ckpt["lr_scheduler"])
        if self.training_dataloader is not None and ckpt.get("data_sampler") is not None and self.curriculum_learning_enabled():
            self.training_dataloader.data_sampler.load_state_dict(ckpt["data_sampler"])
        if hasattr(self.optimizer, "groups"):
            ok = self._load_zero_checkpoint(load_dir, tag, load_optimizer_states=load_optimizer_states)
            if not ok and is_z3:
                return None, None
        elif load_optimizer_states and self.optimizer is not None and ckpt.get("optimizer") is not None:
            self.optimizer.load_state_dict(ckpt["optimizer"])
    elif hasattr(self.optimizer, "refresh_fp32_params"):
        self.optimizer.refresh_fp32_params()
    skip = {"module", "buffer_names", "optimizer", "param_shapes", "frozen_param_shapes", "shared_params",
            "frozen_param_fragments", "lr_scheduler", "data_sampler", "random_ltd", "sparse_tensor_module_names",
            "skipped_steps", "global_steps", "global_samples", "dp_world_size", "mp_world_size", "ds_config",
            "ds_version", "num_experts"}
    client_state = {k: v for k, v in ckpt.items() if k not in skip}
    return path, client_state


def _get_all_zero_checkpoints(self, load_dir, tag):
    bf16 = "bf16_" if self.bfloat16_enabled() else ""
    mp = _mp_rank(self)
    pat = os.path.join(load_dir, str(tag), f"{bf16}zero_pp_rank_*_mp_rank_{mp:02d}_optim_states.pt")
    files = glob.glob(pat)
    if not files and bf16:
        files = glob.glob(os.path.join(load_dir, str(tag), f"zero_pp_rank_*_mp_rank_{mp:02d}_optim_states.pt"))
    import re

    def rank_of(f):
        return int(re.search(r"zero_pp_rank_(\d+)_", os.path.basename(f)).group(1))
    return sorted(files, key=rank_of)


def _load_zero_checkpoint(self, load_dir, tag, load_optimizer_states=True):
    files = self._get_all_zero_checkpoints(load_dir, tag)
    if not files:
        logger.warning(f"no ZeRO optimizer state files found under {os.path.join(load_dir, str(tag))}")
        return False
    opt = self.optimizer
    my_file_rank = _zero_file_rank(self)
    saved_world_total = len(files)
    slice_world = dist.get_world_size(opt.dp_process_group) if dist.is_initialized() else 1
    slice_rank = dist.get_rank(opt.dp_process_group) if dist.is_initialized() else 0
    same_layout = saved_world_total == (self.seq_dp_world_size if dist.is_initialized() else 1)
    if same_layout:
        sd = self.checkpoint_engine.load(files[my_file_rank], map_location="cpu")
        sds = [sd["optimizer_state_dict"]]
    else:
        # different DP size: load every partition (of the first replica) and re-partition
        first = self.checkpoint_engine.load(files[0], map_location="cpu")["optimizer_state_dict"]
        pc = first.get("partition_count")
        saved_slice = pc[0] if isinstance(pc, (list, tuple)) else pc
        # Re-partitioning to a new data-parallel size is always available here: the optimizer files carry unpadded
        # per-group state plus ``param_offsets``.  ``elastic_checkpoint`` is therefore accepted and has no effect
        # (the reference refuses this load unless elastic/universal: engine.py:3238-3243).
        sds = [first] + [self.checkpoint_engine.load(f, map_location="cpu")["optimizer_state_dict"]
                         for f in files[1:saved_slice]]
    opt.load_state_dict(sds, load_optimizer_states=load_optimizer_states,
                        load_from_fp32_weights=self.zero_load_from_fp32_weights(),
                        checkpoint_folder=os.path.join(load_dir, str(tag)))
    log_dist(f"loaded {len(sds)} zero partition file(s) for rank {self.global_rank} (dp layout "
             f"{'unchanged' if same_layout else 'changed'})", ranks=[0])
    return True


def _load_moe_experts(self, load_dir, tag, ckpt):
    from ..moe.layer import MoE
    moe_layer_id = 0
    module_sd = ckpt["module"]
    for name, module in self.module.named_modules():
        if not isinstance(module, MoE):
            continue
        group_name = module.expert_group_name
        num_local = module.num_local_experts
        expp_rank = groups._get_expert_parallel_rank(group_name)
        prefix = f"{name}.deepspeed_moe.experts.deepspeed_experts."
        for local_id in range(num_local):
            gid = expp_rank * num_local + local_id
            esd = self.checkpoint_engine.load(_get_expert_ckpt_name(load_dir, moe_layer_id, gid, tag, self.mpu),
                                              map_location="cpu")
            for k, v in esd.items():
                module_sd[k.replace(f"{prefix}{gid}.", f"{prefix}{local_id}.")] = v
        moe_layer_id += 1


def _zero3_consolidated_16bit_state_dict(self, exclude_frozen_parameters=False):
    """Full 16-bit state dict on rank 0 (gathers one ZeRO-3 unit at a time)."""
    if not self.zero_optimization_partition_weights():
        raise ValueError("this function requires ZeRO-3 mode")
    opt = self.optimizer
    out = OrderedDict() if self.global_rank == 0 else None
    names = {id(p): n for n, p in self.module.named_parameters()}
    for u in opt.units:
        opt._fetch(u)
        if self.global_rank == 0:
            for fg in u.groups:
                for p in fg.params:
                    if exclude_frozen_parameters and not p.requires_grad:
                        continue
                    out[names[id(p)]] = p.detach().cpu().clone()
        opt._release(u)
    if self.global_rank == 0:
        for n, b in self.module.named_buffers():
            out[n] = b.detach().cpu()
    return out


def save_16bit_model(self, save_dir, save_filename="pytorch_model.bin", exclude_frozen_parameters=False):
    path = os.path.join(save_dir, save_filename)
    if self.zero_optimization_partition_weights():
        if not self.zero_gather_16bit_weights_on_model_save():
            logger.info(f"Did not save the model {path} because `stage3_gather_16bit_weights_on_model_save` is False")
            return False
        sd = self._zero3_consolidated_16bit_state_dict(exclude_frozen_parameters)
    else:
        sd = self.module_state_dict(exclude_frozen_parameters=exclude_frozen_parameters)
    tag = f"global_step{self.global_steps}"
    self.checkpoint_engine.create(tag)
    if self.global_rank == 0:
        os.makedirs(save_dir, exist_ok=True)
        logger.info(f"Saving model weights to {path}, tag: {tag}")
        self.checkpoint_engine.save(sd, path)
    self.checkpoint_engine.commit(tag)
    return True
