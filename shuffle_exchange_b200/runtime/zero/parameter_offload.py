"""``DeepSpeedZeRoOffload``: ZeRO-3 parameter sharding + fetch/release hooks WITHOUT a training optimizer
(ZeRO-Inference, frozen sub-models, the hybrid engine's generation phase).

Reference: ``runtime/zero/parameter_offload.py:86-510`` (hooks that call ``fetch_sub_module`` /
``release_sub_module``; persistent-parameter marking :262-280; leaf-module granularity :512-589).  This class
reuses the unit machinery of :class:`DeepSpeedZeroOptimizer_Stage3` with an inert optimizer: parameters are flattened
per unit and sharded across the data-parallel group; every unit is all-gathered just before it runs and released right
after; with ``offload_param.device = cpu`` the shards live in (pinned) host memory and are streamed over PCIe/C2C.
"""
from __future__ import annotations

import torch

from ...utils.logging import logger
from .partitioned_param_coordinator import PartitionedParameterCoordinator
from .stage3 import DeepSpeedZeroOptimizer_Stage3


class _Inert(torch.optim.Optimizer):
    """Holds the parameter groups; never updates anything."""

    def __init__(self, params):
        super().__init__(params, dict(lr=0.0))

    def step(self, closure=None):
        return None


class DeepSpeedZeRoOffload:
    def __init__(self, module, timers=None, ds_config=None, overlap_comm=True, prefetch_bucket_size=50000000,
                 max_reuse_distance=1000000000, max_live_parameters=1000000000, param_persistence_threshold=100000,
                 model_persistence_threshold=None, dp_process_group=None, offload_param_config=None, mpu=None,
                 zero_param_parallel_group=None, zero_quantized_weights=False,
                 zero_quantized_nontrainable_weights=False, zero_module_granularity_threshold=0,
                 log_trace_cache_warnings=False):
        self.module = module
        params = [p for p in module.parameters()]
        frozen = [p for p in params if not p.requires_grad]
        for p in frozen:                         # sharding machinery keys on requires_grad groups
            p.requires_grad_(True)
        self._frozen = frozen
        self._engine = DeepSpeedZeroOptimizer_Stage3(
            module, _Inert(params), timers=timers, ds_config=ds_config, overlap_comm=overlap_comm,
            prefetch_bucket_size=prefetch_bucket_size, max_reuse_distance=max_reuse_distance,
            max_live_parameters=max_live_parameters, param_persistence_threshold=param_persistence_threshold,
            model_persistence_threshold=model_persistence_threshold, dp_process_group=dp_process_group,
            offload_param_config=offload_param_config, mpu=mpu, fused_step=False, verbose=False,
            zero_quantized_weights=zero_quantized_weights,
            zero_quantized_nontrainable_weights=zero_quantized_nontrainable_weights)
        for p in frozen:
            p.requires_grad_(False)
        self.param_coordinator = PartitionedParameterCoordinator(self._engine)
        self.persistent_parameters = [p for u in self._engine.units if u.persistent for fg in u.groups for p in fg.params]
        logger.info(f"ZeRO-Offload (no optimizer): {len(self._engine.units)} unit(s), "
                    f"{len(self.persistent_parameters)} persistent parameter tensor(s)")

    def get_param_coordinator(self):
        return self.param_coordinator

    def partition_all_parameters(self):
        self._engine.release_all()

    def empty_partition_cache(self):
        self.partition_all_parameters()

    def destroy(self):
        self._engine.destroy()

    def _remove_module_hooks(self):
        self._engine.destroy()

    def setup_zero_stage3_hooks(self):
        pass     # installed by the stage-3 machinery at construction

    def mark_persistent_parameters(self, param_threshold, model_threshold):
        return self.persistent_parameters
