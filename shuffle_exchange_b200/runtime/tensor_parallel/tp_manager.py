"""Training-side TP manager (reference: ``runtime/tensor_parallel/tp_manager.py:12-66``): applies AutoTP to the model
with the groups built by ``utils.groups`` and makes ranks of one TP group consume identical random streams
(same dropout masks on replicated activations) and identical data."""
from __future__ import annotations

import torch

from ... import comm as dist
from ...utils import groups
from ...utils.logging import logger
from .config import AUTOTP_MODE, TPTrainingConfig


class TpTrainingManager:
    def __init__(self, model, tp_size: int, dtype):
        self.module = model
        self.config = self._initialize_config(dtype)
        from ...module_inject.auto_tp import tp_model_init
        from ...module_inject.layers import set_autotp_mode
        set_autotp_mode(training=True)
        group = None
        if dist.is_initialized() and tp_size > 1:
            try:
                group = groups.get_tensor_model_parallel_group()
            except Exception:
                group = None
        self.module = tp_model_init(model, tp_size=tp_size, dtype=dtype, tp_group=group)
        self._broadcast_replicated(group)
        self._seed_like_group_leader(group)

    def _initialize_config(self, dtype) -> TPTrainingConfig:
        return TPTrainingConfig(dtype=dtype)

    @torch.no_grad()
    def _broadcast_replicated(self, group):
        """Parameters that AutoTP left replicated (norms, embeddings, biases of row-parallel layers) must start
        identical inside a TP group."""
        if group is None or dist.get_world_size(group=group) <= 1:
            return
        src = dist.get_global_rank(group, 0)
        for p in self.module.parameters():
            if not getattr(p, "tensor_model_parallel", False):
                dist.broadcast(p.data, src=src, group=group)

    def _seed_like_group_leader(self, group):
        if group is None or dist.get_world_size(group=group) <= 1:
            return
        seed = torch.tensor([torch.initial_seed() % (2 ** 31)], dtype=torch.int64,
                            device=torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else "cpu")
        dist.broadcast(seed, src=dist.get_global_rank(group, 0), group=group)
        torch.manual_seed(int(seed.item()))
        logger.info(f"TP rank {dist.get_rank(group=group)}: RNG seeded with the group leader's seed {int(seed.item())}")
