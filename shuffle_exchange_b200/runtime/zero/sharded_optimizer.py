"""ShardedOptimizerBase — mixed-precision optimizer over :class:`FlatGroup` storage.

Everything the reference re-implements separately in ``FP16_Optimizer``, ``BF16_Optimizer``,
``DeepSpeedZeroOptimizer`` (stage 1/2) and ``DeepSpeedZeroOptimizer_Stage3`` — loss scaling and
overflow skipping, global gradient norm + clipping, the partitioned step on fp32 master shards, the
low-precision write-back, gradient accumulation across micro-steps, state-dict plumbing — lives
here once.  Subclasses decide how FlatGroups are formed (per optimizer param group, or per ZeRO-3
parameter unit), when gradients are reduced, and when parameters are (re)gathered.

Hot path on B200: one ``sxb_sumsq`` launch per group for the norm, then one fused Adam launch per
group that reads the bf16 gradient shard, applies ``1/loss_scale x clip`` on the fly, updates the
fp32 master + moments and writes the bf16 shard — no unscale pass, no fp32 gradient copy, no
separate fp32->bf16 cast (the reference does all three: stage_1_and_2.py:2117-2177).
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional

import torch

from ...comm import comm as dist
from ...device import get_accelerator
from ...native import lib as _n
from ...utils.logging import logger
from ..fp16.loss_scaler import CreateLossScaler
from .flat_group import FlatGroup

BASE_OPTIMIZER_STATE = "base_optimizer_state"
BASE_OPTIMIZER_STATE_STEP = "base_optimizer_state_step"
SINGLE_PARTITION_OF_FP32_GROUPS = "single_partition_of_fp32_groups"
GROUP_PADDINGS = "group_paddings"
PARTITION_COUNT = "partition_count"
ZERO_STAGE = "zero_stage"
CLIP_GRAD = "clip_grad"
PARAM_SLICE_MAPPINGS = "param_slice_mappings"
LOSS_SCALER = "loss_scaler"
DS_VERSION = "ds_version"


class ShardedOptimizerBase:
    """Common machinery; not used directly."""

    zero_stage = 0

    def __init__(self, init_optimizer, timers=None, dtype=torch.bfloat16, static_loss_scale=1.0,
                 dynamic_loss_scale=False, dynamic_loss_args=None, clip_grad=0.0, dp_process_group=None, mpu=None,
                 gradient_accumulation_steps=1, gradient_accumulation_dtype=None, communication_data_type=None,
                 gradient_predivide_factor=1.0, postscale_gradients=True, check_grad_overflow=None):
        self.optimizer = init_optimizer
        self.timers = timers
        self.dtype = dtype
        self.clip_grad = float(clip_grad or 0.0)
        self.mpu = mpu
        self.dp_process_group = dp_process_group
        self.gradient_accumulation_steps = gradient_accumulation_steps
        self.gradient_predivide_factor = gradient_predivide_factor
        self.postscale_gradients = postscale_gradients
        self.communication_data_type = communication_data_type or dtype
        self.gradient_accumulation_dtype = gradient_accumulation_dtype or dtype
        self.device = torch.device(get_accelerator().current_device_name())
        self.groups: List[FlatGroup] = []
        self.hp_params: List[torch.nn.Parameter] = []
        self.micro_step_id = 0
        self.overflow = False
        self._global_grad_norm = 0.0
        self.skipped_steps = 0
        self.custom_loss_scaler = False
        self.external_loss_scale = None
        self.loss_scaler = CreateLossScaler(dtype=dtype, static_loss_scale=static_loss_scale,
                                            dynamic_scaling=dynamic_loss_scale, dynamic_loss_args=dynamic_loss_args)
        self.dynamic_loss_scale = self.loss_scaler.dynamic
        self.check_overflow_enabled = (dtype == torch.float16) if check_grad_overflow is None else check_grad_overflow
        self._noop = None     # device int flag: 1 => skip step (overflow)
        self._stats = None    # device float[2]: sumsq, nonfinite count
        self._fused = hasattr(init_optimizer, "step_flat")

    # ---------------------------------------------------------------------- optimizer passthrough
    @property
    def param_groups(self):
        return self.optimizer.param_groups

    @param_groups.setter
    def param_groups(self, v):
        self.optimizer.param_groups = v

    @property
    def state(self):
        return self.optimizer.state

    @property
    def loss_scale(self):
        if self.custom_loss_scaler:
            return self.external_loss_scale
        return self.loss_scaler.cur_scale

    cur_scale = loss_scale

    def override_loss_scale(self, loss_scale):
        if loss_scale != self.external_loss_scale:
            logger.info(f"[sxb200] setting loss scale from {self.external_loss_scale} -> {loss_scale}")
        self.custom_loss_scaler = True
        self.external_loss_scale = loss_scale

    def get_lr(self):
        return self.optimizer.param_groups[0]["lr"]

    def set_lr(self, lr):
        for g in self.optimizer.param_groups:
            g["lr"] = lr

    # ------------------------------------------------------------------------ group registration
    def _register_group(self, fg: FlatGroup, param_group: dict):
        """Make ``fg.hp_shard`` the (only) parameter the base optimizer sees for this group."""
        hp = torch.nn.Parameter(fg.hp_shard, requires_grad=True)
        fg.hp_param = hp
        fg.param_group = param_group
        self.groups.append(fg)
        self.hp_params.append(hp)
        return hp

    def _device_scalars(self):
        if self._noop is None:
            self._noop = torch.zeros(1, dtype=torch.int32, device=self.device)
            self._stats = torch.zeros(2, dtype=torch.float32, device=self.device)
        return self._noop, self._stats

    # ------------------------------------------------------------------------------- backward
    def backward(self, loss, retain_graph=False):
        self.micro_step_id += 1
        self._pre_backward()
        scale = self.loss_scale
        if scale != 1.0:
            (loss.float() * scale).backward(retain_graph=retain_graph)
        else:
            loss.backward(retain_graph=retain_graph)
        self._post_backward()

    def _pre_backward(self):
        for g in self.groups:
            g.attach_main_grads()

    def _post_backward(self):
        for g in self.groups:
            g.collect_autograd_grads()

    # ------------------------------------------------------------------------------ grad norm
    def _shard_sumsq(self) -> torch.Tensor:
        """Device tensor [sumsq, nonfinite] over this rank's gradient shards (own partition only)."""
        _, stats = self._device_scalars()
        stats.zero_()
        for g in self.groups:
            gs = g.grad_shard
            if gs is None:
                continue
            valid = self._valid_shard_numel(g)
            if valid <= 0:
                continue
            if gs.is_cuda:
                _n.call("sxb_sumsq", gs.data_ptr(), valid, _n.dtype_code(gs.dtype), stats.data_ptr(), _n.stream_ptr())
            else:
                f = gs[:valid].float()
                bad = ~torch.isfinite(f)
                stats[0] += torch.where(bad, torch.zeros_like(f), f).pow(2).sum()
                stats[1] += bad.sum()
        return stats

    def _valid_shard_numel(self, g: FlatGroup) -> int:
        """Elements of this rank's partition that are real parameters (excludes tail padding; the
        inter-parameter alignment gaps are always zero)."""
        lo = g.rank * g.part
        return max(0, min(g.part, g.unpadded_numel - lo))

    def _norm_reduce_groups(self):
        """Groups over which partial squared norms must be summed: the sharding group, plus the
        tensor-parallel group when an mpu is present (model-parallel params are disjoint)."""
        out = []
        if self.dp_process_group is not None and dist.get_world_size(self.dp_process_group) > 1:
            out.append(self.dp_process_group)
        if self.mpu is not None:
            for n in ("get_tensor_model_parallel_group", "get_model_parallel_group"):
                if hasattr(self.mpu, n):
                    g = getattr(self.mpu, n)()
                    if g is not None and dist.get_world_size(g) > 1:
                        out.append(g)
                    break
        return out

    def compute_grad_norm_and_overflow(self):
        """Returns (global_norm (python float, already divided by the loss scale), overflow flag)."""
        stats = self._shard_sumsq().clone()
        for grp in self._norm_reduce_groups():
            dist.all_reduce(stats, op=dist.ReduceOp.SUM, group=grp)
        sumsq, bad = stats.tolist()
        overflow = bad > 0 or not math.isfinite(sumsq)
        norm = math.sqrt(max(sumsq, 0.0)) / self.loss_scale if not overflow else float("inf")
        return norm, overflow

    # ------------------------------------------------------------------------------------ step
    def _combined_scale(self, norm: float) -> float:
        """Multiplier applied to raw (scaled) gradients inside the optimizer kernel."""
        s = 1.0 / self.loss_scale
        if self.clip_grad > 0.0 and norm > self.clip_grad:
            s *= self.clip_grad / (norm + 1e-6)
        return s

    def _step_group(self, fg: FlatGroup, grad_scale: float, write_lp: bool = True):
        hp, pg = fg.hp_param, fg.param_group
        lp_out = fg.lp_shard if write_lp else None
        if self._fused:
            grad = fg.grad_shard
            if grad.device != hp.device:
                grad = grad.to(hp.device, non_blocking=False)
            if lp_out is not None and lp_out.device != hp.device:
                # master and low-precision shard live on different devices (optimizer offload, or offload_param with a
                # device-resident master): the kernel writes a staging shard next to the master, then one copy
                stage = getattr(fg, "_lp_stage", None)
                if stage is None or stage.device != hp.device:
                    stage = torch.empty(fg.part, dtype=fg.lp_dtype, device=hp.device,
                                        pin_memory=(hp.device.type == "cpu" and get_accelerator().is_available()))
                    fg._lp_stage = stage
                self.optimizer.step_flat(hp, grad, pg, lp_out=stage, grad_scale=grad_scale)
                lp_out.copy_(stage, non_blocking=True)
            else:
                self.optimizer.step_flat(hp, grad, pg, lp_out=lp_out, grad_scale=grad_scale)
        else:
            g32 = fg.grad_shard.to(device=hp.device, dtype=torch.float32)
            if grad_scale != 1.0:
                g32.mul_(grad_scale)
            hp.grad = g32
            saved = [grp["params"] for grp in self.optimizer.param_groups]
            try:
                for grp in self.optimizer.param_groups:
                    grp["params"] = [hp] if grp is pg else []
                self.optimizer.step()
            finally:
                for grp, ps in zip(self.optimizer.param_groups, saved):
                    grp["params"] = ps
            hp.grad = None
            if lp_out is not None:
                lp_out.copy_(hp.data)
        if lp_out is not None:
            fg.lp_shard_written()         # NVMe-resident shards: the store writes the buffer back before reusing it

    def _update_scale(self, overflow: bool):
        if not self.custom_loss_scaler:
            self.loss_scaler.update_scale(overflow)

    def zero_grad(self, set_to_none=True):
        for g in self.groups:
            for p in g.params:
                p.grad = None

    # ------------------------------------------------------------- reference optimizer-object surface
    # The reference's ZeRO optimizers expose these to the engine, to ``utils/tensor_fragment`` and to user scripts
    # (``stage3.py:2315-2563``, ``stage_1_and_2.py:1995-2056``).  Here they are thin views over the flat groups.
    def has_overflow(self, partition_gradients=True) -> bool:
        """Any inf / nan in this step's (reduced) gradients, agreed across the data-parallel group."""
        return bool(self.compute_grad_norm_and_overflow()[1])

    def check_overflow(self, partition_gradients=True) -> None:
        self.overflow = self.has_overflow(partition_gradients)

    def has_overflow_serial(self, params=None) -> bool:
        ps = params if params is not None else [p for g in self.groups for p in g.params]
        return any(p.grad is not None and not bool(torch.isfinite(p.grad).all()) for p in ps)

    def scaled_global_norm(self, norm_type=2):
        """Global gradient norm of the loss-SCALED gradients (multiply-by-loss-scale of ``get_global_grad_norm``)."""
        norm, overflow = self.compute_grad_norm_and_overflow()
        return float("inf") if overflow else norm * self.loss_scale

    def unscale_and_clip_grads(self, sub_group_id_or_grads=None, total_norm=None) -> float:
        """Factor the step kernels apply to the raw gradients: 1 / loss_scale, times the clip ratio when
        ``total_norm`` (unscaled) exceeds ``clip_grad``.  Gradients are not touched here — unscale and clip are folded
        into the fused optimizer kernel."""
        norm = total_norm if total_norm is not None else self.compute_grad_norm_and_overflow()[0]
        return self._combined_scale(norm)

    def get_param_id(self, param) -> int:
        ids = self.__dict__.setdefault("_param_ids", {})
        if not ids:
            for g in self.groups:
                for p in g.params:
                    ids[id(p)] = len(ids)
        return ids[id(param)]

    def get_bit16_param_group(self, group_no: int):
        return list(self.groups[group_no].params)

    def get_groups_without_padding(self, groups_with_padding=None):
        """fp32 shard of every group with the alignment padding of the last rank removed (checkpoint layout)."""
        out = []
        for g in self.groups:
            n = self._valid_shard_numel(g)
            out.append(g.hp_param.data[:n])
        return out

    def get_lean_optimizer_state(self):
        """Per group, the base optimizer's state tensors without padding (reference ``_get_base_optimizer_state``)."""
        out = []
        for g in self.groups:
            n = self._valid_shard_numel(g)
            st = self._optimizer_state_for(g)
            out.append({k: (v[:n] if torch.is_tensor(v) and v.dim() == 1 and v.numel() >= n else v)
                        for k, v in st.items()})
        return out

    def get_fp32_grad_partitions(self):
        return {i: {0: g.grad_shard} for i, g in enumerate(self.groups) if g.grad_shard is not None}

    def refresh_fp32_params(self):
        """fp32 master shards := current low-precision shards (after weights were loaded into the module)."""
        for g in self.groups:
            if g.lp_shard is not None:
                g.hp_param.data.copy_(g.lp_shard.to(g.hp_param.device, torch.float32))

    def backward_prologue(self):
        self._pre_backward()

    def backward_epilogue(self):
        self._post_backward()

    def checkpoint_event_prologue(self):
        """Before a checkpoint is written: nothing is swapped out in this design unless NVMe offload is active."""
        sw = getattr(self, "optimizer_swapper", None)
        if sw is not None and hasattr(sw, "swap_in_all"):
            sw.swap_in_all()

    def checkpoint_event_epilogue(self):
        sw = getattr(self, "optimizer_swapper", None)
        if sw is not None and hasattr(sw, "swap_out_all"):
            sw.swap_out_all()

    def reset_swap_buffers(self):
        sw = getattr(self, "optimizer_swapper", None)
        if sw is not None and hasattr(sw, "reset_buffers"):
            sw.reset_buffers()

    def dump_pre_step_gradients(self, debug_fp32_grads=None):
        """Debug aid (reference ``stage3.py:2176-2188``): norm of every group's reduced gradient shard."""
        for i, g in enumerate(self.groups):
            if g.grad_shard is not None:
                print(f"Pre-Step Dump Norms for Group {i} grad_norm = {float(g.grad_shard.float().norm()):.6f}")

    def dump_post_step_gradients(self):
        for i, g in enumerate(self.groups):
            print(f"Post-Step Dump Norms for Group {i} fp32_norm = {float(g.hp_param.data.float().norm()):.6f}")

    def print_rank_0(self, message):
        if dist.get_rank() == 0:
            print(message)

    def offload_states(self, include=None, device="cpu", pin_memory=True, non_blocking=False):
        from .offload_states import offload_optimizer_states
        return offload_optimizer_states(self, include=include, device=device, pin_memory=pin_memory,
                                        non_blocking=non_blocking)

    def reload_states(self, non_blocking=False):
        from .offload_states import reload_optimizer_states
        return reload_optimizer_states(self, non_blocking=non_blocking)

    # tensor-fragment API on the optimizer object (the reference attaches the optimizer to every parameter as
    # ``param._z3_optimizer`` and ``utils/tensor_fragment`` calls these)
    def get_full_hp_param(self, param, optim_state_key=None):
        from ...utils import tensor_fragment as tf
        return tf.safe_get_full_fp32_param(param) if optim_state_key is None else \
            tf.safe_get_full_optimizer_state(param, optim_state_key)

    def set_full_hp_param(self, value, param, optim_state_key=None):
        from ...utils import tensor_fragment as tf
        return tf.safe_set_full_fp32_param(param, value) if optim_state_key is None else \
            tf.safe_set_full_optimizer_state(param, value, optim_state_key)

    def get_fp32_grad_for_param(self, param):
        from ...utils import tensor_fragment as tf
        return tf.safe_get_full_grad(param)

    def set_fp32_grad_for_param(self, value, param):
        from ...utils import tensor_fragment as tf
        return tf.safe_set_full_grad(param, value)

    def update_fp32_grad_for_param_vectorized(self, update_func, param_list):
        from ...utils import tensor_fragment as tf
        return tf.safe_update_full_grad_vectorized(param_list, update_func)

    def get_local_fp32_param(self, param, optim_state_key=None):
        from ...utils import tensor_fragment as tf
        return tf.safe_get_local_fp32_param(param) if optim_state_key is None else \
            tf.safe_get_local_optimizer_state(param, optim_state_key)

    def set_local_hp_param(self, value, param, optim_state_key=None):
        from ...utils import tensor_fragment as tf
        return tf.safe_set_local_fp32_param(param, value) if optim_state_key is None else \
            tf.safe_set_local_optimizer_state(param, value, optim_state_key)

    def get_local_fp32_grad_for_param(self, param):
        from ...utils import tensor_fragment as tf
        return tf.safe_get_local_grad(param)

    def set_local_grad_for_param(self, value, param):
        from ...utils import tensor_fragment as tf
        return tf.safe_set_local_grad(param, value)

    def get_hp_param_device(self, param, optim_state_key=None):
        link = None
        try:
            from ...utils.tensor_fragment import get_hp_fragment_mapping
            link = get_hp_fragment_mapping(param)
        except Exception:
            pass
        for g in self.groups:
            if any(p is param for p in g.params):
                if optim_state_key is None:
                    return g.hp_param.device
                st = self._optimizer_state_for(g).get(optim_state_key)
                return st.device if torch.is_tensor(st) else g.hp_param.device
        return None

    # ------------------------------------------------------------------------------ state dict
    def _base_state_dict(self):
        """Base optimizer state with tensors keyed by group index (one fp32 shard per group)."""
        sd = self.optimizer.state_dict()
        return sd

    def _optimizer_state_for(self, fg: FlatGroup) -> Dict[str, torch.Tensor]:
        return self.optimizer.state.get(fg.hp_param, {})
