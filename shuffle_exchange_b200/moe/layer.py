"""``MoE`` — the user-facing mixture-of-experts layer (reference: deepspeed/moe/layer.py:18-132).

``forward`` returns ``(output, l_aux, exp_counts)``; the expert-parallel groups ``ep_size_{n}`` are
created lazily by :meth:`set_deepspeed_parallelism` (called from the engine).  Sub-module names
(``deepspeed_moe.gate.wg``, ``deepspeed_moe.experts.deepspeed_experts.{i}``) are part of the checkpoint
format and are kept."""
from typing import Optional

import torch
from torch import nn
from torch.nn import functional as F

from ..comm import comm as dist
from ..utils import groups
from ..utils.logging import log_dist
from .experts import Experts
from .sharded_moe import MOELayer, TopKGate


class MoE(nn.Module):
    def __init__(self, hidden_size: int, expert: nn.Module, num_experts: int = 1, ep_size: int = 1, k: int = 1,
                 capacity_factor: float = 1.0, eval_capacity_factor: float = 1.0, min_capacity: int = 4,
                 use_residual: bool = False, noisy_gate_policy: Optional[str] = None, drop_tokens: bool = True,
                 use_rts: bool = True, use_tutel: bool = False, enable_expert_tensor_parallelism: bool = False,
                 top2_2nd_expert_sampling: bool = True) -> None:
        super().__init__()
        self.use_residual = use_residual
        self.enable_expert_tensor_parallelism = enable_expert_tensor_parallelism
        assert num_experts % ep_size == 0, f"Number of experts ({num_experts}) should be divisible by expert parallel size ({ep_size})"
        self.ep_size = ep_size
        self.expert_group_name = f"ep_size_{self.ep_size}"
        self.num_experts = num_experts
        self.num_local_experts = num_experts // self.ep_size
        log_dist(f"Creating MoE layer with num_experts: {num_experts} | num_local_experts: {self.num_local_experts} | "
                 f"expert_parallel_size: {self.ep_size}", [0])
        assert noisy_gate_policy is None or noisy_gate_policy in ["None", "Jitter", "RSample"], \
            "Unsupported noisy_gate_policy: " + str(noisy_gate_policy)
        experts = Experts(expert, self.num_local_experts, self.expert_group_name)
        self.deepspeed_moe = MOELayer(
            TopKGate(hidden_size, num_experts, k, capacity_factor, eval_capacity_factor, min_capacity,
                     noisy_gate_policy, drop_tokens, use_rts, None, top2_2nd_expert_sampling),
            experts, self.expert_group_name, self.ep_size, self.num_local_experts, use_tutel=use_tutel)
        if self.use_residual:
            self.mlp = expert
            self.coefficient = nn.Linear(hidden_size, 2)     # residual-MoE mixing coefficients

    def set_deepspeed_parallelism(self, use_data_before_expert_parallel_: bool = False) -> None:
        self._create_process_groups(use_data_before_expert_parallel_=use_data_before_expert_parallel_)

    def _create_process_groups(self, use_data_before_expert_parallel_: bool = False) -> None:
        if not dist.is_initialized():
            return
        if self.expert_group_name not in groups._get_expert_parallel_group_dict():
            if groups.mpu is None or not self.enable_expert_tensor_parallelism:
                groups._create_expert_and_data_parallel(self.ep_size, use_data_before_expert_parallel_)
            else:
                groups._create_expert_data_and_model_parallel(self.ep_size, groups.mpu, use_data_before_expert_parallel_)
        self.deepspeed_moe._set_ep_group(groups._get_expert_parallel_group(self.expert_group_name))

    def forward(self, hidden_states: torch.Tensor, used_token: Optional[torch.Tensor] = None):
        output = self.deepspeed_moe(hidden_states, used_token)
        if self.use_residual:
            output_mlp = self.mlp(hidden_states)
            if isinstance(output_mlp, tuple):
                output_mlp = output_mlp[0]
            coef = F.softmax(self.coefficient(hidden_states), dim=-1)
            output = output * coef[..., 0:1] + output_mlp * coef[..., 1:]
        return output, self.deepspeed_moe.l_aux, self.deepspeed_moe.exp_counts
