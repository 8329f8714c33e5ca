"""MiCS: ZeRO-3 sharded over a SUB-group of the data-parallel world, replicated across sub-groups.

Reference: ``runtime/zero/mics.py:58-472`` (``MiCS_Init`` :58, hierarchical all-gather :254-321, ``MiCS_Optimizer``
:349-472 with the cross-replica gradient all-reduce :432-460).  Here both classes are thin configurations of the
unit-based ZeRO-3 implementation: parameters and optimizer state are sharded over ``param_shard_group`` (so the
all-gathers stay inside e.g. one NVLink domain), the reduce-scatter runs in that group, and before the optimizer
step every rank all-reduces its gradient shard with the ranks holding the same shard in the other replicas.
"""
from __future__ import annotations

import torch

from ... import comm as dist
from ...utils.logging import log_dist
from .mics_utils import MiCS_CommGroups, create_mics_comm_groups
from .partition_parameters import Init
from .stage3 import DeepSpeedZeroOptimizer_Stage3


def has_hierarchical_all_gather_groups(comm_groups: MiCS_CommGroups) -> bool:
    return comm_groups.param_intra_node_group is not None and comm_groups.param_inter_node_shard_group is not None


def mics_hierarchical_all_gather(shard: torch.Tensor, out: torch.Tensor, comm_groups: MiCS_CommGroups) -> torch.Tensor:
    """Two-hop all-gather of equally sized shards: first across nodes (few large messages over the slow links), then
    inside the node.  ``out`` receives the shards in shard-rank order (node-major)."""
    inter, intra = comm_groups.param_inter_node_shard_group, comm_groups.param_intra_node_group
    n_inter, n_intra = dist.get_world_size(inter), dist.get_world_size(intra)
    part = shard.numel()
    stage = torch.empty(n_inter * part, dtype=shard.dtype, device=shard.device)
    dist.all_gather_into_tensor(stage, shard.contiguous().view(-1), group=inter)
    tmp = torch.empty(n_intra * n_inter * part, dtype=shard.dtype, device=shard.device)
    dist.all_gather_into_tensor(tmp, stage, group=intra)
    # tmp is [local][node][part]; shard-rank order is [node][local][part]
    out.view(n_inter, n_intra, part).copy_(tmp.view(n_intra, n_inter, part).transpose(0, 1))
    return out


class MiCS_Init(Init):
    """``zero.Init`` whose partitioning group is the MiCS shard group taken from
    ``zero_optimization.mics_shard_size`` (reference :58-140)."""

    def __init__(self, module=None, data_parallel_group=None, sequence_data_parallel_group=None,
                 mem_efficient_linear=True, remote_device=None, pin_memory=False, config_dict_or_path=None, config=None,
                 enabled=True, dtype=None, mpu=None):
        assert config_dict_or_path is not None or config is not None, "Must provide configuration for MiCS Initialization"
        from ..config import DeepSpeedConfig
        cfg = config_dict_or_path if config_dict_or_path is not None else config
        ds_config = cfg if isinstance(cfg, DeepSpeedConfig) else DeepSpeedConfig(cfg, mpu)
        if not dist.is_initialized():
            dist.init_distributed()
        zc = ds_config.zero_config
        self.mics_comm_groups = create_mics_comm_groups(zc.mics_shard_size, data_parallel_group,
                                                        hierarchical_allgather=zc.mics_hierarchical_params_gather,
                                                        mpu=mpu)
        super().__init__(module=module, data_parallel_group=self.mics_comm_groups.param_shard_group,
                         mem_efficient_linear=mem_efficient_linear, remote_device=remote_device, pin_memory=pin_memory,
                         config_dict_or_path=cfg, enabled=enabled, dtype=dtype, mpu=mpu)

    def _convert_module(self, module):
        super()._convert_module(module)
        for p in module.parameters():
            p.comm = self.mics_comm_groups


class MiCS_Optimizer(DeepSpeedZeroOptimizer_Stage3):
    def __init__(self, module, init_optimizer, ds_config=None, dp_process_group=None, mics_shard_size=-1,
                 mics_hierarchical_params_gather=False, mpu=None, **kwargs):
        log_dist("Init MiCS optimizer", ranks=[0])
        first = next(p for g in init_optimizer.param_groups for p in g["params"])
        groups = getattr(first, "comm", None)
        if not isinstance(groups, MiCS_CommGroups):
            if mics_shard_size is None or mics_shard_size <= 0:
                zc = getattr(ds_config, "zero_config", None)
                mics_shard_size = getattr(zc, "mics_shard_size", -1) if zc is not None else -1
            assert mics_shard_size > 0, "mics_shard_size must be set for MiCS"
            groups = create_mics_comm_groups(mics_shard_size, dp_process_group,
                                             hierarchical_allgather=mics_hierarchical_params_gather, mpu=mpu)
        self.mics_comm_groups = groups
        kwargs["fused_step"] = False      # the step must wait for the cross-replica reduction
        super().__init__(module, init_optimizer, ds_config=ds_config, dp_process_group=groups.param_shard_group,
                         mpu=mpu, **kwargs)

    def allreduce_mics_shard_grads(self):
        """Average every pending gradient shard over the replicate group (reference :432-460)."""
        g = self.mics_comm_groups
        if g.param_repli_size is None or g.param_repli_size <= 1:
            return
        shards = [fg.grad_shard for fg in self.groups if fg.grad_shard is not None]
        if not shards:
            return
        flat = torch.cat([s.reshape(-1) for s in shards]) if len(shards) > 1 else shards[0].reshape(-1)
        flat.div_(g.param_repli_size)
        dist.all_reduce(flat, group=g.param_repli_group)
        if len(shards) > 1:
            off = 0
            for s in shards:
                s.copy_(flat[off: off + s.numel()].view_as(s))
                off += s.numel()

    def step(self, closure=None):
        self.allreduce_mics_shard_grads()
        return super().step(closure)

    def load_state_dict(self, state_dict_list, load_optimizer_states=True, load_from_fp32_weights=False,
                        checkpoint_folder=None, load_serial=None, param_shapes=None):
        """Checkpoints are written per shard-group rank; every replica loads the same shard files."""
        return super().load_state_dict(state_dict_list, load_optimizer_states=load_optimizer_states,
                                       load_from_fp32_weights=load_from_fp32_weights,
                                       checkpoint_folder=checkpoint_folder, load_serial=load_serial,
                                       param_shapes=param_shapes)
