"""BF16_Optimizer — bf16 parameters with fp32 master partitions and fp32 gradient accumulation
(reference: deepspeed/runtime/bf16_optimizer.py:35-560; selected by the engine for bf16 without
ZeRO, engine.py:1384-1386).  It is the ZeRO-1 machinery over the full data-parallel group with an
fp32 accumulator, so it is expressed as a configuration of :class:`DeepSpeedZeroOptimizer` rather
than a second implementation."""
import torch

from .zero.stage_1_and_2 import DeepSpeedZeroOptimizer


class BF16_Optimizer(DeepSpeedZeroOptimizer):
    def __init__(self, init_optimizer, param_names=None, mpu=None, clip_grad=0.0, norm_type=2,
                 allgather_bucket_size=5000000000, dp_process_group=None, timers=None, grad_acc_dtype=None,
                 graph_harvesting=False, immediate_grad_update=True, has_moe_layers=False, check_grad_overflow=False,
                 gradient_accumulation_steps=1):
        assert grad_acc_dtype in (None, torch.float32, torch.bfloat16), \
            f"BF16Optimizer: Unsupported gradient accumulation data type: {grad_acc_dtype}"
        super().__init__(init_optimizer, param_names=param_names, timers=timers, static_loss_scale=1.0,
                         dynamic_loss_scale=False, allgather_bucket_size=allgather_bucket_size,
                         dp_process_group=dp_process_group, mpu=mpu, clip_grad=clip_grad,
                         gradient_accumulation_dtype=grad_acc_dtype or torch.float32,
                         communication_data_type=torch.bfloat16, gradient_accumulation_steps=gradient_accumulation_steps,
                         partition_grads=False, has_moe_layers=has_moe_layers, check_grad_overflow=check_grad_overflow,
                         dtype=torch.bfloat16, replica_averaging=False, verbose=False)
        self.zero_stage = 1
        self.immediate_grad_update = immediate_grad_update
        self.norm_type = norm_type
        self.bf16_groups = self.bit16_groups
        self.fp32_groups_flat_partition = self.single_partition_of_fp32_groups

    def update_hp_grads(self, clear_lp_grads=False):
        """Pipeline engine hook: fold freshly produced lp grads into the accumulator."""
        for g in self.groups:
            g.collect_autograd_grads()

    def clear_lp_grads(self):
        self.zero_grad()

    def get_grads_for_reduction(self):
        return [g.flat_grad for g in self.groups if g.flat_grad is not None], {}

    # ---- reference-named gradient plumbing (``bf16_optimizer.py:330-470``)
    def create_grad_acc_hooks(self):
        """Gradient hand-off hooks are installed when the flat groups are built (main-grad views the wgrad kernels
        write into, plus post-accumulate hooks for everything else); nothing to add afterwards."""
        return None

    def accumulate_hp_grads_and_remove_lp(self, lp_param, group_idx, param_idx):
        """Fold ``lp_param.grad`` into the group's high-precision accumulator and drop the low-precision gradient."""
        g = self.groups[group_idx]
        if lp_param.grad is not None:
            g.ensure_flat_grad()
            view = g.grad_view(param_idx)
            view.add_(lp_param.grad.reshape(view.shape).to(view.dtype))
            lp_param.grad = None

    def clear_hp_grads(self):
        for g in self.groups:
            if g.flat_grad is not None:
                g.flat_grad.zero_()
            if g.grad_shard is not None and (g.flat_grad is None or g.grad_shard.data_ptr() != g.flat_grad.data_ptr()):
                g.grad_shard.zero_()

    def get_grads_for_norm(self, for_clipping=False):
        """Gradient tensors that enter the global norm on this rank.  ``for_clipping``: also return them grouped by
        flat group (the reference returns ``(non_expert_grads, expert_grads)``; experts here are separate groups whose
        norm is reduced over their own process group by ``compute_grad_norm_and_overflow``)."""
        grads = [g.grad_shard if g.grad_shard is not None else g.flat_grad for g in self.groups]
        grads = [t for t in grads if t is not None]
        if not for_clipping:
            return grads
        experts = {}
        for g in self.groups:
            name = (g.param_group or {}).get("name") if getattr(g, "param_group", None) else None
            if name and (g.param_group or {}).get("moe", False):
                t = g.grad_shard if g.grad_shard is not None else g.flat_grad
                if t is not None:
                    experts.setdefault(name, []).append(t)
        flat_experts = {id(t) for ts in experts.values() for t in ts}
        return [t for t in grads if id(t) not in flat_experts], experts
