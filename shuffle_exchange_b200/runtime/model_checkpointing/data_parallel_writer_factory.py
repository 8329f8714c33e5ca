"""Who writes which slice of a replicated checkpoint file.

Reference: ``runtime/model_checkpointing/data_parallel_writer_factory.py:11-216``.  Ranks of one data-parallel (or
expert-data-parallel) group hold identical model states; instead of rank 0 writing the whole file, the file is cut into
``world_size`` byte ranges written concurrently.  The parallel unit picks the writers:

* ``None``     — classic: only dp-rank 0 writes everything;
* ``REPLICA``  — every replica writes ``1/dp`` of the file;
* ``SOCKET`` / ``MACHINE`` — one writer per CPU socket / per node among the replicas (NVMe and PCIe bandwidth is per
  socket/node, more writers than that only add contention).

Formulation used here: the replicas of my model slice are grouped by the resource (socket or machine) they sit on, the
lowest rank of every group is that resource's writer, and writers are numbered in resource order."""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional

from ..zero.config import ZeroStageEnum
from .constants import CheckpointDataParallel


@dataclass
class DataParallelWriterConfig:
    world_size: int
    rank: int
    global_rank: int
    local_rank: int
    pure_dp: bool


class DataParallelWriterFactory:

    def __init__(self, uni_parallel_info, parallel_unit):
        self._info = uni_parallel_info
        self._parallel_unit = parallel_unit
        self._num_resources = uni_parallel_info.num_sockets if parallel_unit == CheckpointDataParallel.SOCKET \
            else uni_parallel_info.num_machines
        self._ranks_per_resource = max(1, uni_parallel_info.global_world_size // max(1, self._num_resources))

    # ------------------------------------------------------------------------------------------ public
    def create_config(self, zero_stage, has_moe_layers) -> Optional[DataParallelWriterConfig]:
        if zero_stage == ZeroStageEnum.weights:
            return self._create_config(1, 0)          # ZeRO-3: every rank owns unique shards and writes its own file
        cfg = self._expert_config() if has_moe_layers else self._dense_config()
        if cfg is None and zero_stage >= ZeroStageEnum.optimizer_states:
            return self._create_config(1, 0)          # optimizer shards are unique per rank
        return cfg

    # ------------------------------------------------------------------------------------------ helpers
    def _create_config(self, world_size, rank) -> DataParallelWriterConfig:
        return DataParallelWriterConfig(world_size=world_size, rank=rank, global_rank=self._info.global_rank,
                                        local_rank=self._info.local_rank, pure_dp=self._info.pure_dp)

    def _writers_among(self, replica_ranks: List[int]) -> Optional[DataParallelWriterConfig]:
        """One writer per resource among ``replica_ranks`` (global ranks holding the same data as this rank)."""
        by_resource = {}
        for r in sorted(replica_ranks):
            by_resource.setdefault(r // self._ranks_per_resource, []).append(r)
        writers = [ranks[0] for _, ranks in sorted(by_resource.items())]
        me = self._info.global_rank
        if me not in writers:
            return None
        return self._create_config(len(writers), writers.index(me))

    def _dense_config(self):
        mpu = self._info.mpu_info
        if self._parallel_unit is None:
            dp_rank = self._info.global_rank if mpu is None else mpu.dp_rank
            return self._create_config(1, 0) if dp_rank == 0 else None
        if self._info.pure_dp:
            return self._ddp(self._info.global_world_size, self._info.global_rank,
                             list(range(self._info.global_world_size)))
        if self._parallel_unit == CheckpointDataParallel.REPLICA:
            return self._create_config(mpu.dp_world_size, mpu.dp_rank)
        return self._writers_among(list(mpu.dp_peer_ranks))

    def _expert_config(self):
        ep = self._info.ep_info
        if self._parallel_unit is None:
            return self._create_config(1, 0) if ep.dp_rank == 0 else None
        assert self._info.pure_dp, "3D parallelism is not yet supported for data parallel checkpointing."
        if self._parallel_unit == CheckpointDataParallel.REPLICA or ep.ep_world_size == 1:
            return self._ddp(ep.dp_world_size, ep.dp_rank, list(ep.dp_peer_ranks))
        return self._writers_among(list(ep.dp_peer_ranks))

    def _ddp(self, dp_world_size, dp_rank, peer_ranks):
        if self._parallel_unit == CheckpointDataParallel.REPLICA:
            return self._create_config(dp_world_size, dp_rank)
        n_machines = self._info.num_machines
        if self._parallel_unit == CheckpointDataParallel.SOCKET:
            if dp_world_size == n_machines:                     # one rank per machine: all of them write
                return self._create_config(n_machines, dp_rank)
            n_sockets = self._info.num_sockets
            per_socket = max(1, dp_world_size // n_sockets)
            if dp_rank % per_socket == 0:
                return self._create_config(n_sockets, dp_rank // per_socket)
            return None
        per_machine = max(1, dp_world_size // n_machines)
        if dp_rank % per_machine == 0:
            return self._create_config(n_machines, self._info.machine_rank)
        return None
