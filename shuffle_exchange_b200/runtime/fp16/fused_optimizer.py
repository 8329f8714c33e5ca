"""FP16_Optimizer — fp16 parameters, flat fp32 master copy, dynamic loss scaling, no sharding
(reference: deepspeed/runtime/fp16/fused_optimizer.py).  Gradients are averaged over the data-parallel
group by the optimizer itself (one all-reduce per flat group) and the master is replicated, i.e. the
ZeRO-1 machinery with a partition world of 1."""
import torch

from ...comm import comm as dist
from ..zero.flat_group import FlatGroup
from ..zero.stage_1_and_2 import DeepSpeedZeroOptimizer


class FP16_Optimizer(DeepSpeedZeroOptimizer):
    def __init__(self, init_optimizer, deepspeed=None, static_loss_scale=1.0, dynamic_loss_scale=False,
                 initial_dynamic_scale=2 ** 32, dynamic_loss_args=None, verbose=True, mpu=None, clip_grad=0.0,
                 fused_adam_legacy=False, has_moe_layers=False, timers=None, dp_process_group=None):
        self._dp_group_for_allreduce = dp_process_group
        first = next(p for g in init_optimizer.param_groups for p in g["params"])
        # world-of-one sharding: every rank owns the whole master; DP averaging is done in _reduce
        super().__init__(init_optimizer, timers=timers, static_loss_scale=static_loss_scale,
                         dynamic_loss_scale=dynamic_loss_scale, dynamic_loss_args=dynamic_loss_args, mpu=mpu,
                         clip_grad=clip_grad, dp_process_group=None, partition_grads=False, shard_optimizer=False,
                         has_moe_layers=has_moe_layers, check_grad_overflow=(first.dtype == torch.float16),
                         dtype=first.dtype, replica_averaging=False, verbose=False,
                         gradient_accumulation_dtype=first.dtype)
        self.zero_stage = 0
        self.fp16_groups = self.bit16_groups
        self.fp32_groups_flat = self.single_partition_of_fp32_groups

    def _launch_group_reduce(self, fg: FlatGroup):
        if getattr(fg, "_reduced_this_micro_step", False):
            return
        fg._reduced_this_micro_step = True
        grp = self._dp_group_for_allreduce
        if dist.is_initialized() and dist.get_world_size(grp) > 1:
            fg.flat_grad.div_(dist.get_world_size(grp))
            dist.all_reduce(fg.flat_grad, group=grp)
        fg.grad_shard = fg.flat_grad[: fg.part]

    def _norm_reduce_groups(self):
        out = []
        if self.mpu is not None and hasattr(self.mpu, "get_model_parallel_group"):
            g = self.mpu.get_model_parallel_group()
            if g is not None and dist.get_world_size(g) > 1:
                out.append(g)
        return out
