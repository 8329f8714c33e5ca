"""Common base for the framework's optimizer wrappers (reference: deepspeed/runtime/base_optimizer.py:14-81:
``DeepSpeedOptimizer`` marker class and ``ZeROOptimizer`` with ``load_hp_checkpoint_state_from_checkpoint_dir``,
which walks the universal-checkpoint folder and loads per-parameter fp32 fragments)."""
from __future__ import annotations

import os

import torch

from ..utils.logging import logger


class DeepSpeedOptimizer:
    pass


class ZeROOptimizer(DeepSpeedOptimizer):

    def load_hp_checkpoint_state_from_checkpoint_dir(self, lp_groups_name: str, checkpoint_dir: str) -> None:
        """Universal-checkpoint load: ``<dir>/zero/<param name>/{fp32,exp_avg,exp_avg_sq}.pt`` hold FULL tensors;
        every rank cuts out the slice that belongs to its shard."""
        from ..checkpoint.universal_checkpoint import load_hp_checkpoint_state
        zero_dir = os.path.join(checkpoint_dir, "zero")
        optim_sd_path = os.path.join(checkpoint_dir, "optimizer_state.pt")
        if os.path.isfile(optim_sd_path):
            optim_sd = torch.load(optim_sd_path, map_location="cpu", weights_only=False)
            self._load_global_state(optim_sd)
        tp_rank = 0
        tp_world = 1
        mpu = getattr(self, "mpu", None)
        if mpu is not None:
            from ..utils.bwc import bwc_tensor_model_parallel_rank, bwc_tensor_model_parallel_world_size
            tp_rank, tp_world = bwc_tensor_model_parallel_rank(mpu), bwc_tensor_model_parallel_world_size(mpu)
        steps = []
        for name, lp in self._named_lp_params():
            folder = os.path.join(zero_dir, name)
            if not os.path.isdir(folder):
                logger.warning(f"universal checkpoint has no entry for {name}")
                continue
            step = load_hp_checkpoint_state(self, name, lp, folder, tp_rank, tp_world)
            if step is not None:
                steps.append(int(step))
        if steps:
            self._set_optimizer_step(max(steps))

    # hooks the sharded optimizers implement ---------------------------------------------------
    def _load_global_state(self, sd):  # pragma: no cover - overridden
        pass

    def _named_lp_params(self):  # pragma: no cover - overridden
        return []

    def _set_optimizer_step(self, step):  # pragma: no cover - overridden
        pass
