"""Local expert container (reference: deepspeed/moe/experts.py:13-38): ``num_local_experts`` deep
copies of the user's expert module; parameters are tagged ``allreduce=False`` and ``group_name`` so
the engine reduces them over the expert-data-parallel group only."""
import copy
from typing import List, Optional

import torch
from torch import nn


class Experts(nn.Module):
    def __init__(self, expert: nn.Module, num_local_experts: int = 1, expert_group_name: Optional[str] = None):
        super().__init__()
        self.deepspeed_experts = nn.ModuleList([copy.deepcopy(expert) for _ in range(num_local_experts)])
        self.num_local_experts = num_local_experts
        for e in self.deepspeed_experts:
            for p in e.parameters():
                p.allreduce = False
                p.group_name = expert_group_name

    def forward(self, inputs: torch.Tensor) -> torch.Tensor:
        """inputs [ep, local_E, C, M] -> same shape; expert i processes inputs[:, i]."""
        chunks = inputs.chunk(self.num_local_experts, dim=1)
        outs: List[torch.Tensor] = []
        for chunk, expert in zip(chunks, self.deepspeed_experts):
            out = expert(chunk)
            if isinstance(out, tuple):
                out = out[0]        # ignore a bias term returned alongside
            outs.append(out)
        return torch.cat(outs, dim=1)
